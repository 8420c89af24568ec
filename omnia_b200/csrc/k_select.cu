// Candidate selection, exact rescoring and final ordering.
//
// The scorers (k_score_stream.cu, k_score_umma.cu) leave two packed bf16-accurate
// keys per (query, sub-bucket).  This file turns them into the reference's answer:
//   select_keys   the K2 best keys of a query            (radix select, exact)
//   rescore       exact cosine distance of each survivor  (fp32 inputs, fp64 math:
//                 pgvector's cosine_distance formula, the `<=>` the reference calls at
//                 internal/memory/store_read.go:268,278,279)
//   finalize      ORDER BY distance LIMIT k with the `<= maxDistance` cutoff
//                 (store_read.go:260,278-280) plus the exactness certificate
//   exact_scan /  the fallback for queries whose certificate fails: every eligible row
//   select_exact  is scored exactly, then the k nearest are selected
//   merge_topk    G shards' local answers -> the global one
#include "kernels.cuh"

namespace omr {

namespace {

constexpr int kSelThreads = 256;

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ double clamp_sim(double s) { return s > 1.0 ? 1.0 : (s < -1.0 ? -1.0 : s); }

// Finds the `need`-th largest value among keys that match (u & mask) == prefix, one radix
// digit per call.  Thread 0 walks the histogram; everyone gets the result through smem.
struct RadixState {
  uint32_t hist[256];
  unsigned long long prefix;
  unsigned long long mask;
  uint32_t need;
  uint32_t digit;
};

__device__ __forceinline__ void radix_pick_desc(RadixState* rs, int shift) {
  // largest-first: walk digits from 255 down
  if (threadIdx.x == 0) {
    uint32_t cum = 0, need = rs->need;
    int d = 255;
    for (; d > 0; d--) {
      if (cum + rs->hist[d] >= need) break;
      cum += rs->hist[d];
    }
    rs->need = need - cum;
    rs->prefix |= (unsigned long long)d << shift;
    rs->mask |= 0xFFull << shift;
  }
}
__device__ __forceinline__ void radix_pick_asc(RadixState* rs, int shift) {
  if (threadIdx.x == 0) {
    uint32_t cum = 0, need = rs->need;
    int d = 0;
    for (; d < 255; d++) {
      if (cum + rs->hist[d] >= need) break;
      cum += rs->hist[d];
    }
    rs->need = need - cum;
    rs->prefix |= (unsigned long long)d << shift;
    rs->mask |= 0xFFull << shift;
  }
}

// One block per query.  Output: sel_rows[q][K2] (shard-local row, -1 = empty) and
// sel_bound[q] = an upper bound on the approximate score of any key NOT selected
// (-inf when every valid key was selected).
__global__ void __launch_bounds__(kSelThreads) select_keys_kernel(const int32_t* __restrict__ cand, int64_t kpq, int S,
                                                                  int K2, int K2x, int32_t* __restrict__ sel_rows,
                                                                  float* __restrict__ sel_bound) {
  __shared__ RadixState rs;
  __shared__ uint32_t s_valid, s_out, s_eq;
  const int q = blockIdx.x;
  const int32_t* keys = cand + (int64_t)q * kpq;
  int32_t* out = sel_rows + (int64_t)q * K2x;
  if (threadIdx.x == 0) {
    s_valid = 0;
    s_out = 0;
    s_eq = 0;
  }
  for (int i = threadIdx.x; i < K2; i += kSelThreads) out[i] = -1;
  __syncthreads();
  uint32_t local = 0;
  for (int64_t i = threadIdx.x; i < kpq; i += kSelThreads) local += keys[i] != kKeyNone;
  if (local) atomicAdd(&s_valid, local);
  __syncthreads();
  const uint32_t valid = s_valid;
  if (valid <= (uint32_t)K2) {
    for (int64_t i = threadIdx.x; i < kpq; i += kSelThreads) {
      const int32_t k = keys[i];
      if (k != kKeyNone) out[atomicAdd(&s_out, 1u)] = (int32_t)((i >> 1) * S + (k & 0xFF));
    }
    if (threadIdx.x == 0) sel_bound[q] = -INFINITY;
    return;
  }
  if (threadIdx.x == 0) {
    rs.prefix = 0;
    rs.mask = 0;
    rs.need = (uint32_t)K2;
  }
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256; i += kSelThreads) rs.hist[i] = 0;
    __syncthreads();
    const uint32_t prefix = (uint32_t)rs.prefix, mask = (uint32_t)rs.mask;
    for (int64_t i = threadIdx.x; i < kpq; i += kSelThreads) {
      const uint32_t u = (uint32_t)keys[i] ^ 0x80000000u;  // kKeyNone -> 0, the smallest
      if ((u & mask) == prefix) atomicAdd(&rs.hist[(u >> shift) & 255u], 1u);
    }
    __syncthreads();
    radix_pick_desc(&rs, shift);
    __syncthreads();
  }
  const uint32_t T = (uint32_t)rs.prefix, need_eq = rs.need;
  for (int64_t i = threadIdx.x; i < kpq; i += kSelThreads) {
    const int32_t k = keys[i];
    const uint32_t u = (uint32_t)k ^ 0x80000000u;
    bool take = u > T;
    if (u == T) take = atomicAdd(&s_eq, 1u) < need_eq;
    if (take) out[atomicAdd(&s_out, 1u)] = (int32_t)((i >> 1) * S + (k & 0xFF));
  }
  if (threadIdx.x == 0) sel_bound[q] = key_score_upper((int32_t)(T ^ 0x80000000u));
}

// One warp per candidate slot.  fp32 products are exact in fp64, so the result is the
// mathematically exact cosine up to fp64 rounding of a <=2000-term sum.  Two modes:
//   dense  (items == nullptr): slots [0, K2) of every query;
//   list   (items != nullptr): the flat slot indices finalize pass 1 appended for
//          suspect sub-buckets; the count lives on the device, so the grid is fixed and
//          warps stride over the list.
__global__ void __launch_bounds__(256) rescore_kernel(ShardView sh, const float* __restrict__ q,
                                                      const double* __restrict__ qnorm2, int nq, int K2, int K2x,
                                                      const int32_t* __restrict__ items,
                                                      const int32_t* __restrict__ item_count,
                                                      const int32_t* __restrict__ sel_rows,
                                                      double* __restrict__ sel_dist) {
  const int lane = threadIdx.x & 31;
  const int64_t warps_total = (int64_t)gridDim.x * 8;
  const int64_t total = items ? (int64_t)*item_count : (int64_t)nq * K2;
  for (int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); w < total; w += warps_total) {
    int64_t flat;
    if (items) flat = items[w];
    else flat = (w / K2) * K2x + (w % K2);
    const int qi = (int)(flat / K2x);
    const int32_t row = sel_rows[flat];
    if (row < 0) {
      if (lane == 0) sel_dist[flat] = INFINITY;
      continue;
    }
    const float* m = sh.rows_f32 + (int64_t)row * sh.pitch;
    const float* qv = q + (int64_t)qi * sh.dim;
    double dot = 0.0;
    for (int i = lane; i < sh.dim; i += 32) dot += (double)qv[i] * (double)m[i];
    dot = warp_sum_d(dot);
    if (lane == 0) {
      const double d = 1.0 - clamp_sim(dot / sqrt(qnorm2[qi] * sh.norm2[row]));
      sel_dist[flat] = (d == d) ? d : INFINITY;  // NaN (zero-norm query) never ranks
    }
  }
}

struct SortItem {
  double d;
  int32_t row;
};
__device__ __forceinline__ bool item_less(const SortItem& a, const SortItem& b) {
  return a.d < b.d || (a.d == b.d && a.row < b.row);
}

__device__ void bitonic_sort(double* sd, int32_t* sr, int P) {
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < (P >> 1); i += blockDim.x) {
        const int lo = 2 * i - (i & (stride - 1));
        const int hi = lo + stride;
        const bool up = ((lo & size) == 0);
        SortItem a{sd[lo], sr[lo]}, b{sd[hi], sr[hi]};
        if (item_less(b, a) == up) {
          sd[lo] = b.d; sr[lo] = b.row;
          sd[hi] = a.d; sr[hi] = a.row;
        }
      }
    }
  }
  __syncthreads();
}

// Pass 1, one block per query: ORDER BY distance over the K2 rescored candidates, the
// `<= maxDistance` cutoff, LIMIT k -- then the exactness certificate.
//
// Certificate.  A row that was not rescored has an approximate score no larger than
// (a) sel_bound[q] if its key lost the K2 selection, or (b) the second key of its
// sub-bucket if it never got a key.  It can only belong in the answer if its exact
// similarity reaches tau (the k-th best found, or the cutoff), i.e. if its approximate
// score reaches thr = tau - margin.  So: (a) >= thr  -> hard fail (exact fallback);
// (b) >= thr -> the sub-bucket is "suspect" and ALL its eligible rows are appended to the
// query's extra slots for pass 2.  Adding candidates only raises tau, so pass 2 needs no
// second certificate.
__global__ void __launch_bounds__(256) finalize_kernel(
    int K2, int K2x, int P, int k, int S, int32_t* __restrict__ sel_rows, const double* __restrict__ sel_dist,
    const double* __restrict__ max_dist, const float* __restrict__ sel_bound, const int32_t* __restrict__ cand,
    int64_t kpq, float margin, const uint32_t* __restrict__ bitmaps, int words_per_group,
    const int32_t* __restrict__ query_group, int64_t row_base, int64_t row_stride,
    int64_t* __restrict__ out_rows, double* __restrict__ out_dist, int32_t* __restrict__ out_count,
    int32_t* __restrict__ extra_cnt,
    int32_t* __restrict__ extra_items, int32_t* __restrict__ extra_total, int32_t* __restrict__ fail_list,
    int32_t* __restrict__ fail_count, const int32_t* __restrict__ qlist) {
  extern __shared__ __align__(16) uint8_t fsm[];
  double* sd = reinterpret_cast<double*>(fsm);
  int32_t* sr = reinterpret_cast<int32_t*>(sd + P);
  __shared__ int s_fail, s_extra;
  const int q = qlist ? qlist[blockIdx.x] : blockIdx.x;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    const int32_t r = i < K2 ? sel_rows[(int64_t)q * K2x + i] : -1;
    sd[i] = r >= 0 ? sel_dist[(int64_t)q * K2x + i] : INFINITY;
    sr[i] = r >= 0 ? r : INT32_MAX;
  }
  if (threadIdx.x == 0) {
    s_fail = 0;
    s_extra = 0;
  }
  bitonic_sort(sd, sr, P);
  const double maxd = max_dist ? max_dist[q] : INFINITY;
  int found;
  {  // entries within the cutoff form a prefix of the sorted list
    int lo = 0, hi = P < k ? P : k;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (sd[mid] <= maxd && sd[mid] < INFINITY) lo = mid + 1; else hi = mid;
    }
    found = lo;
  }
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    const bool ok = i < found;
    out_rows[(int64_t)q * k + i] = ok ? row_base + (int64_t)sr[i] * row_stride : -1;
    out_dist[(int64_t)q * k + i] = ok ? sd[i] : INFINITY;
  }
  if (threadIdx.x == 0) out_count[q] = found;
  if (cand == nullptr) return;

  double tau;
  if (found >= k) tau = 1.0 - sd[k - 1];
  else if (maxd < INFINITY) tau = 1.0 - maxd;
  else tau = -INFINITY;
  const float thr = (float)(tau - (double)margin) - 1e-7f;
  if (threadIdx.x == 0 && sel_bound[q] >= thr) s_fail = 1;
  const int32_t* keys = cand + (int64_t)q * kpq;
  const int E = K2x - K2;
  const uint32_t* bm = bitmaps + (int64_t)(query_group ? query_group[q] : 0) * words_per_group;
  const int64_t nsb = kpq >> 1;
  for (int64_t sb = threadIdx.x; sb < nsb; sb += blockDim.x) {
    const int32_t k2 = keys[2 * sb + 1];
    if (k2 == kKeyNone || key_score_upper(k2) < thr) continue;
    for (int w = 0; w < S / 32; w++) {
      uint32_t bits = bm[sb * (S / 32) + w];
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const int slot = atomicAdd(&s_extra, 1);
        if (slot < E) {
          const int32_t flat = (int32_t)((int64_t)q * K2x + K2 + slot);
          sel_rows[flat] = (int32_t)(sb * S + w * 32 + b);
          extra_items[atomicAdd(extra_total, 1)] = flat;
        } else {
          s_fail = 1;
        }
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    extra_cnt[q] = s_fail ? 0 : (s_extra < E ? s_extra : E);
    if (s_fail) fail_list[atomicAdd(fail_count, 1)] = q;
  }
}

// Pass 2, one block per query; only queries with extra candidates do anything.  Sorts the
// K2 + extra rescored candidates, drops the duplicates (a suspect sub-bucket's two keyed
// rows are in both segments), and rewrites the answer.
__global__ void __launch_bounds__(256) finalize_extra_kernel(int K2, int K2x, int P, int k,
                                                             const int32_t* __restrict__ sel_rows,
                                                             const double* __restrict__ sel_dist,
                                                             const double* __restrict__ max_dist,
                                                             const int32_t* __restrict__ extra_cnt, int64_t row_base,
                                                             int64_t row_stride, int64_t* __restrict__ out_rows,
                                                             double* __restrict__ out_dist,
                                                             int32_t* __restrict__ out_count) {
  extern __shared__ __align__(16) uint8_t fsm[];
  double* sd = reinterpret_cast<double*>(fsm);
  int32_t* sr = reinterpret_cast<int32_t*>(sd + P);
  const int q = blockIdx.x;
  const int ne = extra_cnt[q];
  if (ne <= 0) return;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    const int32_t r = i < K2 + ne ? sel_rows[(int64_t)q * K2x + i] : -1;
    sd[i] = r >= 0 ? sel_dist[(int64_t)q * K2x + i] : INFINITY;
    sr[i] = r >= 0 ? r : INT32_MAX;
  }
  bitonic_sort(sd, sr, P);
  if (threadIdx.x == 0) {
    const double maxd = max_dist ? max_dist[q] : INFINITY;
    int found = 0;
    int32_t prev = -1;
    for (int i = 0; i < P && found < k; i++) {
      if (!(sd[i] <= maxd) || !(sd[i] < INFINITY)) break;
      if (sr[i] == prev) continue;
      prev = sr[i];
      out_rows[(int64_t)q * k + found] = row_base + (int64_t)sr[i] * row_stride;
      out_dist[(int64_t)q * k + found] = sd[i];
      found++;
    }
    out_count[q] = found;
    for (int i = found; i < k; i++) {
      out_rows[(int64_t)q * k + i] = -1;
      out_dist[(int64_t)q * k + i] = INFINITY;
    }
  }
}

// Exact fallback scan.  grid = (row chunks of 256, slots); warp w of a block owns 32 rows.
__global__ void __launch_bounds__(256) exact_scan_kernel(ShardView sh, const float* __restrict__ q,
                                                         const double* __restrict__ qnorm2,
                                                         const int32_t* __restrict__ slot_query,
                                                         const uint32_t* __restrict__ bitmaps, int words_per_group,
                                                         const int32_t* __restrict__ query_group,
                                                         double* __restrict__ dist_out, int64_t n_pad) {
  const int slot = blockIdx.y;
  const int qi = slot_query[slot];
  const int g = query_group ? query_group[qi] : 0;
  const int lane = threadIdx.x & 31;
  const int64_t word = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int64_t row0 = word * 32;
  if (row0 >= n_pad) return;
  uint32_t bits = bitmaps[(int64_t)g * words_per_group + word];
  double* out = dist_out + (int64_t)slot * n_pad;
  if (!((bits >> lane) & 1u)) out[row0 + lane] = INFINITY;
  const float* qv = q + (int64_t)qi * sh.dim;
  const double qn = qnorm2[qi];
  while (bits) {
    const int b = __ffs(bits) - 1;
    bits &= bits - 1;
    const int64_t row = row0 + b;
    // same summation order as the rescore of the candidate paths (k_post.cu exact_distance2: lane-strided
    // float4, x y z w in turn, xor-shuffle reduction), so a query gets the same bits whichever path answers it
    const float4* m4 = reinterpret_cast<const float4*>(sh.rows_f32 + row * sh.pitch);
    double dot = 0.0;
    for (int i = lane; i < (sh.pitch >> 2); i += 32) {
      const float4 a = __ldg(m4 + i);
      const int e = 4 * i;
      dot = fma((double)a.x, e < sh.dim ? (double)qv[e] : 0.0, dot);
      dot = fma((double)a.y, e + 1 < sh.dim ? (double)qv[e + 1] : 0.0, dot);
      dot = fma((double)a.z, e + 2 < sh.dim ? (double)qv[e + 2] : 0.0, dot);
      dot = fma((double)a.w, e + 3 < sh.dim ? (double)qv[e + 3] : 0.0, dot);
    }
    dot = warp_sum_d(dot);
    if (lane == 0) {
      const double d = 1.0 - clamp_sim(dot / sqrt(qn * sh.norm2[row]));
      out[row] = (d == d) ? d : INFINITY;
    }
  }
}

// One block per slot: the k smallest of n_pad exact distances (all >= 0 or +inf, so their
// bit patterns order like unsigned integers), written into the query's candidate slots.
__global__ void __launch_bounds__(1024) select_exact_kernel(const double* __restrict__ dist, int64_t n_pad,
                                                            const int32_t* __restrict__ slot_query, int K2, int K2x, int k,
                                                            int32_t* __restrict__ sel_rows,
                                                            double* __restrict__ sel_dist) {
  __shared__ RadixState rs;
  __shared__ uint32_t s_out;
  __shared__ unsigned long long s_finite;
  const int slot = blockIdx.x;
  const int q = slot_query[slot];
  const unsigned long long* d = reinterpret_cast<const unsigned long long*>(dist + (int64_t)slot * n_pad);
  int32_t* orow = sel_rows + (int64_t)q * K2x;
  double* odist = sel_dist + (int64_t)q * K2x;
  const unsigned long long kInf = 0x7FF0000000000000ull;
  if (threadIdx.x == 0) {
    s_out = 0;
    s_finite = 0;
    rs.prefix = 0;
    rs.mask = 0;
  }
  for (int i = threadIdx.x; i < K2; i += blockDim.x) {
    orow[i] = -1;
    odist[i] = INFINITY;
  }
  __syncthreads();
  unsigned long long fin = 0;
  for (int64_t i = threadIdx.x; i < n_pad; i += blockDim.x) fin += d[i] < kInf;
  if (fin) atomicAdd(&s_finite, fin);
  __syncthreads();
  const unsigned long long finite = s_finite;
  if (finite == 0) return;
  unsigned long long T = kInf - 1;
  uint32_t need_eq = 0xffffffffu;
  if (finite > (unsigned long long)k) {
    if (threadIdx.x == 0) rs.need = (uint32_t)k;
    for (int shift = 56; shift >= 0; shift -= 8) {
      for (int i = threadIdx.x; i < 256; i += blockDim.x) rs.hist[i] = 0;
      __syncthreads();
      const unsigned long long prefix = rs.prefix, mask = rs.mask;
      for (int64_t i = threadIdx.x; i < n_pad; i += blockDim.x) {
        const unsigned long long u = d[i];
        if ((u & mask) == prefix) atomicAdd(&rs.hist[(u >> shift) & 255ull], 1u);
      }
      __syncthreads();
      radix_pick_asc(&rs, shift);
      __syncthreads();
    }
    T = rs.prefix;
    need_eq = rs.need;
  }
  // Rows strictly nearer than T are all taken; of the rows AT distance T exactly `need_eq` are, and
  // they are the ones with the smallest row ids (an ordered block scan), so the answer does not
  // depend on thread timing even when exact duplicates straddle the k-th place.
  __shared__ uint32_t s_warp_ties[32];
  __shared__ uint32_t s_tie_base;
  if (threadIdx.x == 0) s_tie_base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (int64_t i0 = 0; i0 < n_pad; i0 += blockDim.x) {
    const int64_t i = i0 + threadIdx.x;
    const unsigned long long u = i < n_pad ? d[i] : kInf;
    const bool finite_row = u < kInf;
    const bool tie = finite_row && u == T && finite > (unsigned long long)k;
    const uint32_t ballot = __ballot_sync(0xffffffffu, tie);
    if (lane == 0) s_warp_ties[wid] = __popc(ballot);
    __syncthreads();
    uint32_t before = s_tie_base + __popc(ballot & ((1u << lane) - 1u));
    for (int w = 0; w < wid; w++) before += s_warp_ties[w];
    bool take = finite_row && (finite <= (unsigned long long)k || u < T || (tie && before < need_eq));
    if (take) {
      const uint32_t o = atomicAdd(&s_out, 1u);
      if (o < (uint32_t)K2) {
        orow[o] = (int32_t)i;
        odist[o] = __longlong_as_double((long long)u);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      uint32_t t = 0;
      for (int w = 0; w < nwarps; w++) t += s_warp_ties[w];
      s_tie_base += t;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(256) merge_topk_kernel(const int64_t* __restrict__ in_rows,
                                                         const double* __restrict__ in_dist, int G, int nq, int k,
                                                         int64_t shard_stride, int P,
                                                         int64_t* __restrict__ out_rows,
                                                         double* __restrict__ out_dist,
                                                         int32_t* __restrict__ out_count) {
  extern __shared__ __align__(16) uint8_t msm[];
  double* sd = reinterpret_cast<double*>(msm);
  int64_t* sr = reinterpret_cast<int64_t*>(sd + P);
  const int q = blockIdx.x;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    int64_t r = -1;
    double d = INFINITY;
    if (i < G * k) {
      const int g = i / k, j = i % k;
      r = in_rows[(int64_t)g * shard_stride + (int64_t)q * k + j];
      d = in_dist[(int64_t)g * shard_stride + (int64_t)q * k + j];
    }
    sd[i] = r >= 0 ? d : INFINITY;
    sr[i] = r >= 0 ? r : INT64_MAX;
  }
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < (P >> 1); i += blockDim.x) {
        const int lo = 2 * i - (i & (stride - 1));
        const int hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const double da = sd[lo], db = sd[hi];
        const int64_t ra = sr[lo], rb = sr[hi];
        const bool b_less = db < da || (db == da && rb < ra);
        if (b_less == up) {
          sd[lo] = db; sr[lo] = rb;
          sd[hi] = da; sr[hi] = ra;
        }
      }
    }
  }
  __syncthreads();
  int cnt = 0;
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    const bool ok = i < P && sr[i] != INT64_MAX;
    out_rows[(int64_t)q * k + i] = ok ? sr[i] : -1;
    out_dist[(int64_t)q * k + i] = ok ? sd[i] : INFINITY;
  }
  if (threadIdx.x == 0) {
    const int lim = P < k ? P : k;
    while (cnt < lim && sr[cnt] != INT64_MAX) cnt++;
    out_count[q] = cnt;
  }
}

__global__ void __launch_bounds__(256) gather_rows_kernel(const __nv_bfloat16* __restrict__ src, int pitch,
                                                          const int32_t* __restrict__ idx, int64_t m,
                                                          __nv_bfloat16* __restrict__ dst) {
  const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= m) return;
  const uint4* s = reinterpret_cast<const uint4*>(src + (int64_t)idx[r] * pitch);
  uint4* d = reinterpret_cast<uint4*>(dst + r * pitch);
  for (int i = lane; i < pitch / 8; i += 32) d[i] = s[i];
}

// dst row r = src row idx[r], rows of `units` 8-byte words (compaction of the resident matrices and columns)
__global__ void __launch_bounds__(256) gather_units_kernel(const uint2* __restrict__ src, int64_t units,
                                                           const int32_t* __restrict__ idx, int64_t m,
                                                           uint2* __restrict__ dst) {
  const int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= m) return;
  const uint2* s = src + (int64_t)idx[r] * units;
  uint2* d = dst + r * units;
  for (int64_t i = lane; i < units; i += 32) d[i] = s[i];
}

__global__ void iota_kernel(int32_t* p, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = i;
}

__global__ void fill_empty_kernel(int64_t* rows, double* dist, int32_t* count, int nq, int k) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < (int64_t)nq * k) {
    rows[i] = -1;
    dist[i] = INFINITY;
  }
  if (i < nq) count[i] = 0;
}

__global__ void pair_remap_kernel(PairCand* pairs, unsigned long long n, const int32_t* __restrict__ remap) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    PairCand p = pairs[i];
    p.a = remap[p.a];
    p.b = remap[p.b];
    pairs[i] = p;
  }
}

int pow2ceil(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace

int launch_iota(int32_t* p, int n, cudaStream_t st) {
  if (n <= 0) return 0;
  iota_kernel<<<(n + 255) / 256, 256, 0, st>>>(p, n);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_fill_empty(int64_t* rows, double* dist, int32_t* count, int nq, int k, cudaStream_t st) {
  const int64_t t = (int64_t)nq * k;
  fill_empty_kernel<<<(unsigned)((t + 255) / 256), 256, 0, st>>>(rows, dist, count, nq, k);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_pair_remap(PairCand* pairs, unsigned long long n, const int32_t* remap, cudaStream_t st) {
  if (n == 0) return 0;
  pair_remap_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(pairs, n, remap);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_select_keys(const int32_t* cand_keys, int nq, int64_t keys_per_query, int S, int K2, int K2x,
                       int32_t* sel_rows, float* sel_bound, cudaStream_t st) {
  if (nq <= 0) return 0;
  select_keys_kernel<<<nq, kSelThreads, 0, st>>>(cand_keys, keys_per_query, S, K2, K2x, sel_rows, sel_bound);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_rescore(const ShardView& sh, const float* q, const double* qnorm2, int nq, int K2, int K2x,
                   const int32_t* sel_rows, double* sel_dist, cudaStream_t st) {
  const int64_t warps = (int64_t)nq * K2;
  if (warps <= 0) return 0;
  rescore_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(sh, q, qnorm2, nq, K2, K2x, nullptr, nullptr, sel_rows,
                                                             sel_dist);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_rescore_items(const ShardView& sh, const float* q, const double* qnorm2, int K2x, const int32_t* items,
                         const int32_t* item_count, const int32_t* sel_rows, double* sel_dist, int num_sms,
                         cudaStream_t st) {
  rescore_kernel<<<num_sms * 4, 256, 0, st>>>(sh, q, qnorm2, 0, 1, K2x, items, item_count, sel_rows, sel_dist);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_finalize(const FinalizeArgs& a, cudaStream_t st) {
  const int nblocks = a.qlist ? a.nlist : a.nq;
  if (nblocks <= 0) return 0;
  const int P = pow2ceil(a.K2);
  const size_t smem = (size_t)P * 12;
  if (smem > 32 * 1024)  // static shared memory counts against the 48 KB default too
    OMR_CUDA_OK(cudaFuncSetAttribute(finalize_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  finalize_kernel<<<nblocks, 256, smem, st>>>(a.K2, a.K2x, P, a.k, a.S, a.sel_rows, a.sel_dist, a.max_dist,
                                              a.sel_bound, a.cand_keys, a.keys_per_query, a.margin, a.bitmaps,
                                              a.words_per_group, a.query_group, a.row_base, a.row_stride, a.out_rows, a.out_dist,
                                              a.out_count, a.extra_cnt, a.extra_items, a.extra_total, a.fail_list,
                                              a.fail_count, a.qlist);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_finalize_extra(const FinalizeArgs& a, cudaStream_t st) {
  if (a.nq <= 0) return 0;
  const int P = pow2ceil(a.K2x);
  const size_t smem = (size_t)P * 12;
  if (smem > 32 * 1024)  // static shared memory counts against the 48 KB default too
    OMR_CUDA_OK(cudaFuncSetAttribute(finalize_extra_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  finalize_extra_kernel<<<a.nq, 256, smem, st>>>(a.K2, a.K2x, P, a.k, a.sel_rows, a.sel_dist, a.max_dist, a.extra_cnt,
                                                 a.row_base, a.row_stride, a.out_rows, a.out_dist, a.out_count);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_exact_scan(const ShardView& sh, const float* q, const double* qnorm2, const int32_t* slot_query,
                      int nslots, const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                      double* dist_out, cudaStream_t st) {
  if (nslots <= 0 || words_per_group <= 0) return 0;
  dim3 grid((unsigned)((words_per_group + 7) / 8), (unsigned)nslots);
  exact_scan_kernel<<<grid, 256, 0, st>>>(sh, q, qnorm2, slot_query, bitmaps, words_per_group, query_group, dist_out,
                                          (int64_t)words_per_group * 32);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_select_exact(const double* dist, int64_t n, const int32_t* slot_query, int nslots, int K2, int K2x, int k,
                        int32_t* sel_rows, double* sel_dist, cudaStream_t st) {
  if (nslots <= 0) return 0;
  select_exact_kernel<<<nslots, 1024, 0, st>>>(dist, n, slot_query, K2, K2x, k, sel_rows, sel_dist);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_merge_topk(const int64_t* in_rows, const double* in_dist, int G, int nq, int k, int64_t shard_stride,
                      int64_t* out_rows, double* out_dist, int32_t* out_count, cudaStream_t st) {
  if (nq <= 0) return 0;
  const int P = pow2ceil(G * k);
  const size_t smem = (size_t)P * 16;
  if (smem > 200 * 1024) return fail_msg(-2, "merge of %d shards x k=%d exceeds the 12800-entry limit", G, k);
  if (smem > 32 * 1024)  // static shared memory counts against the 48 KB default too
    OMR_CUDA_OK(cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  if (shard_stride <= 0) shard_stride = (int64_t)nq * k;
  static bool carveout_set = false;  // see k_post.cu
  if (!carveout_set) {
    OMR_CUDA_OK(cudaFuncSetAttribute(merge_topk_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    carveout_set = true;
  }
  merge_topk_kernel<<<nq, 256, smem, st>>>(in_rows, in_dist, G, nq, k, shard_stride, P, out_rows, out_dist, out_count);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_gather_row_bytes(const void* src, size_t row_bytes, const int32_t* idx, int64_t m, void* dst, cudaStream_t st) {
  if (m <= 0) return 0;
  if (row_bytes % 8) return fail_msg(-1, "gather: row of %zu bytes is not a multiple of 8", row_bytes);
  gather_units_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(reinterpret_cast<const uint2*>(src), (int64_t)(row_bytes / 8), idx,
                                                              m, reinterpret_cast<uint2*>(dst));
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_gather_rows(const ShardView& sh, const int32_t* idx, int64_t m, __nv_bfloat16* dst, cudaStream_t st) {
  if (m <= 0) return 0;
  gather_rows_kernel<<<(unsigned)((m + 7) / 8), 256, 0, st>>>(sh.rows_bf16, sh.pitch, idx, m, dst);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
