// Fused per-query post-processing: candidate selection, exact rescore, ORDER BY / LIMIT /
// cutoff, exactness certificate and suspect-sub-bucket rescue in ONE kernel, one block per
// query.  It replaces the select_keys -> rescore -> finalize -> rescore(items) ->
// finalize_extra chain of k_select.cu (kept for K2 > 256 and as the exact fallback's tail)
// and removes every intermediate global buffer between the scorer and the answer.
//
// Reference semantics implemented here: `ORDER BY o.embedding <=> $1 LIMIT $6` with the
// `<= $5` cutoff (internal/memory/store_read.go:266-280), distances computed with pgvector's
// cosine_distance formula in fp64 from the fp32 master rows.
//
// Selection without histograms (packed keys of one query crowd into a handful of exponent
// bins, which serialises shared-memory atomics): every thread keeps the maximum of its own
// strided slice of keys; the K2-th largest of those 512 maxima, M, is a lower bound of the
// true K2-th largest key, so "all keys >= M" is a superset of the top K2 and, for keys in
// random order, only a few entries larger.  All of them are rescored; every key that is not
// collected is < M, which is the bound the certificate uses.
#include "kernels.cuh"
#include "ptx.cuh"

namespace omr {

namespace {

constexpr int kPostThreads = 256;
constexpr int kCandCap = 384;     // collected keys per query (>= K2, K2 <= 256)
constexpr int kSuspectCap = 32;   // suspect sub-buckets per query

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Fixed part of the block's shared memory; the fp64 query copy and the (distance, row) sort arrays
// follow it and are sized at launch (sort_cap = candidates + rescued rows that can be merged in-block).
struct PostSmem {
  int32_t lmax[kPostThreads];
  int32_t ckey[kCandCap];
  int32_t cidx[kCandCap];
  int32_t suspects[kSuspectCap];
  int32_t n_cand, n_suspect, n_extra, fail;
};

__device__ __forceinline__ bool item_less(double da, int32_t ra, double db, int32_t rb) {
  return da < db || (da == db && ra < rb);
}

// ascending bitonic sort of (sd, sr)[0, P), P a power of two
__device__ void sort_items(double* sd, int32_t* sr, int P) {
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < (P >> 1); i += kPostThreads) {
        const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const double da = sd[lo], db = sd[hi];
        const int32_t ra = sr[lo], rb = sr[hi];
        if (item_less(db, rb, da, ra) == up) {
          sd[lo] = db; sr[lo] = rb;
          sd[hi] = da; sr[hi] = ra;
        }
      }
    }
  }
  __syncthreads();
}

// exact cosine distance of one row to the block's query (fp64 copy in shared memory); whole warp.
// Loads are issued six 16-byte requests at a time per lane before any math, so one row of 1536
// floats costs two HBM round trips instead of twelve.
__device__ __forceinline__ double exact_distance(const ShardView& sh, int64_t row, const double* qd, double qn2,
                                                 int lane) {
  const float4* m4 = reinterpret_cast<const float4*>(sh.rows_f32 + row * sh.pitch);
  double acc = 0.0;
  const int n4 = sh.pitch >> 2;  // a multiple of 16
  for (int base = 0; base < n4; base += 32 * 6) {
    float4 a[6];
#pragma unroll
    for (int j = 0; j < 6; j++) {
      const int i = base + j * 32 + lane;
      a[j] = i < n4 ? __ldg(m4 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 6; j++) {
      const int i = base + j * 32 + lane;
      if (i < n4) {
        const double* b = qd + 4 * i;
        acc = fma((double)a[j].x, b[0], acc);
        acc = fma((double)a[j].y, b[1], acc);
        acc = fma((double)a[j].z, b[2], acc);
        acc = fma((double)a[j].w, b[3], acc);
      }
    }
  }
  acc = warp_sum_d(acc);
  double s = acc / sqrt(qn2 * sh.norm2[row]);
  s = s > 1.0 ? 1.0 : (s < -1.0 ? -1.0 : s);
  const double d = 1.0 - s;
  return (d == d) ? d : INFINITY;  // NaN (zero-norm query) never ranks
}

// descending bitonic sort of one value per lane, shuffles only
__device__ __forceinline__ int32_t warp_sort_desc(int32_t v, int lane) {
#pragma unroll
  for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int32_t o = __shfl_xor_sync(0xffffffffu, v, stride);
      const bool lower = (lane & stride) == 0;           // this lane keeps the "first" element of the pair
      const bool desc = (lane & size) == 0;              // direction of this bitonic block
      const bool take_max = lower == desc;
      v = take_max ? max(v, o) : min(v, o);
    }
  }
  return v;
}

// ascending (distance, row) sort of one item per lane, shuffles only
__device__ __forceinline__ void warp_sort_items(double& d, int32_t& r, int lane) {
#pragma unroll
  for (int size = 2; size <= 32; size <<= 1) {
#pragma unroll
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const double od = __shfl_xor_sync(0xffffffffu, d, stride);
      const int32_t orr = __shfl_xor_sync(0xffffffffu, r, stride);
      const bool lower = (lane & stride) == 0;
      const bool asc = (lane & size) == 0;
      const bool other_less = od < d || (od == d && orr < r);
      const bool take_other = (lower == asc) ? other_less : !other_less && !(od == d && orr == r);
      if (take_other) {
        d = od;
        r = orr;
      }
    }
  }
}

// Exact distances of up to two rows at once: twelve 16-byte loads in flight per lane.
__device__ __forceinline__ void exact_distance2(const ShardView& sh, int64_t row0, int64_t row1, bool two,
                                                const double* qd, double qn2, int lane, double& d0, double& d1) {
  const float4* m0 = reinterpret_cast<const float4*>(sh.rows_f32 + row0 * sh.pitch);
  const float4* m1 = reinterpret_cast<const float4*>(sh.rows_f32 + row1 * sh.pitch);
  double a0 = 0.0, a1 = 0.0;
  const int n4 = sh.pitch >> 2;
  for (int base = 0; base < n4; base += 32 * 6) {
    float4 x[6], y[6];
#pragma unroll
    for (int j = 0; j < 6; j++) {
      const int i = base + j * 32 + lane;
      const bool in = i < n4;
      x[j] = in ? __ldg(m0 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
      y[j] = (in && two) ? __ldg(m1 + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < 6; j++) {
      const int i = base + j * 32 + lane;
      if (i < n4) {
        const double* b = qd + 4 * i;
        const double b0 = b[0], b1 = b[1], b2 = b[2], b3 = b[3];
        a0 = fma((double)x[j].x, b0, a0);
        a0 = fma((double)x[j].y, b1, a0);
        a0 = fma((double)x[j].z, b2, a0);
        a0 = fma((double)x[j].w, b3, a0);
        a1 = fma((double)y[j].x, b0, a1);
        a1 = fma((double)y[j].y, b1, a1);
        a1 = fma((double)y[j].z, b2, a1);
        a1 = fma((double)y[j].w, b3, a1);
      }
    }
  }
  a0 = warp_sum_d(a0);
  a1 = warp_sum_d(a1);
  auto fin = [&](double acc, int64_t row) {
    double s = acc / sqrt(qn2 * sh.norm2[row]);
    s = s > 1.0 ? 1.0 : (s < -1.0 ? -1.0 : s);
    const double d = 1.0 - s;
    return (d == d) ? d : INFINITY;
  };
  d0 = fin(a0, row0);
  d1 = two ? fin(a1, row1) : INFINITY;
}

__global__ void __launch_bounds__(kPostThreads, 2) post_kernel(PostArgs a) {
  extern __shared__ __align__(16) uint8_t raw[];
  double* qd = reinterpret_cast<double*>(raw);                       // [pitch]
  double* sd_base = qd + a.sh.pitch;                                 // [sort_cap]
  int32_t* sr_base = reinterpret_cast<int32_t*>(sd_base + a.sort_cap);   // [sort_cap]
  PostSmem& sm = *reinterpret_cast<PostSmem*>(sr_base + a.sort_cap);
  const int kSortCap = a.sort_cap;
  const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  auto stamp = [&](int phase) {
    if (a.timing && tid == 0) a.timing[(int64_t)q * 8 + phase] = clock64();
  };
  stamp(0);
  const int32_t* keys = a.cand_keys + (int64_t)q * a.keys_per_query;
  const int64_t kpq = a.keys_per_query;

  for (int i = tid; i < a.sh.pitch; i += kPostThreads)
    qd[i] = i < a.sh.dim ? (double)a.q[(int64_t)q * a.sh.dim + i] : 0.0;
  if (tid == 0) {
    sm.n_cand = 0;
    sm.n_suspect = 0;
    sm.n_extra = 0;
    sm.fail = 0;
  }
  // ---- pass 1: per-thread maximum over a strided slice (int4 = 4 keys per load).  A query's key
  // row starts 8-byte aligned; `head` scalar keys bring the vector body to 16 bytes.
  const int head = (int)min((int64_t)((4 - (((int64_t)q * kpq) & 3)) & 3), kpq);
  const int4* k4 = reinterpret_cast<const int4*>(keys + head);
  const int64_t n4 = (kpq - head) >> 2;
  const int64_t tail0 = head + (n4 << 2);
  int32_t lm = kKeyNone;
  for (int64_t i0 = tid; i0 < n4; i0 += 8 * kPostThreads) {
    int4 v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int64_t i = i0 + (int64_t)j * kPostThreads;
      v[j] = i < n4 ? __ldg(k4 + i) : make_int4(kKeyNone, kKeyNone, kKeyNone, kKeyNone);
    }
#pragma unroll
    for (int j = 0; j < 8; j++) lm = max(max(lm, v[j].x), max(v[j].y, max(v[j].z, v[j].w)));
  }
  if (tid < head) lm = max(lm, keys[tid]);
  for (int64_t i = tail0 + tid; i < kpq; i += kPostThreads) lm = max(lm, keys[i]);
  stamp(1);  // pass 1 done
  if (a.K2 <= 64) {
    // each warp sorts its 32 maxima with shuffles; warp 0 then pops the K2 largest heads of the
    // 16 sorted lists (one list per lane) to find the K2-th largest overall
    sm.lmax[tid] = warp_sort_desc(lm, lane);
    __syncthreads();
    if (warp == 0) {
      int pos = 0;                                  // lanes 0..15: cursor into list `lane`
      int32_t headv = lane < kPostThreads / 32 ? sm.lmax[lane * 32] : kKeyNone;
      int32_t kth = kKeyNone;
      for (int t = 0; t < a.K2; t++) {
        int32_t best = headv;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
        kth = best;
        const unsigned who = __ballot_sync(0xffffffffu, headv == best && lane < kPostThreads / 32);
        if (best == kKeyNone) break;
        if (lane == __ffs(who) - 1) {
          pos++;
          headv = pos < 32 ? sm.lmax[lane * 32 + pos] : kKeyNone;
        }
      }
      if (lane == 0) sm.lmax[0] = kth;
    }
    __syncthreads();
  } else {
    sm.lmax[tid] = lm;
    // descending bitonic sort of the 512 maxima
    for (int size = 2; size <= kPostThreads; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        __syncthreads();
        if (tid < (kPostThreads >> 1)) {
          const int lo = 2 * tid - (tid & (stride - 1)), hi = lo + stride;
          const bool up = ((lo & size) == 0);
          const int32_t x = sm.lmax[lo], y = sm.lmax[hi];
          if ((y > x) == up) {
            sm.lmax[lo] = y;
            sm.lmax[hi] = x;
          }
        }
      }
    }
    __syncthreads();
    if (tid == 0) sm.lmax[0] = sm.lmax[a.K2 - 1];
    __syncthreads();
  }
  stamp(2);  // maxima sorted
  int32_t M = sm.lmax[0];
  if (M == kKeyNone) M = kKeyNone + 1;  // fewer than K2 threads saw an eligible row: take every valid key
  // ---- pass 2: collect every key >= M
  auto collect = [&](int32_t kv, int64_t idx) {
    if (kv >= M) {
      const int pos = atomicAdd(&sm.n_cand, 1);
      if (pos < kCandCap) {
        sm.ckey[pos] = kv;
        sm.cidx[pos] = (int32_t)idx;
      }
    }
  };
  for (int64_t i0 = tid; i0 < n4; i0 += 8 * kPostThreads) {
    int4 v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int64_t i = i0 + (int64_t)j * kPostThreads;
      v[j] = i < n4 ? __ldg(k4 + i) : make_int4(kKeyNone, kKeyNone, kKeyNone, kKeyNone);
    }
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int64_t i = i0 + (int64_t)j * kPostThreads;
      collect(v[j].x, head + 4 * i);
      collect(v[j].y, head + 4 * i + 1);
      collect(v[j].z, head + 4 * i + 2);
      collect(v[j].w, head + 4 * i + 3);
    }
  }
  if (tid < head) collect(keys[tid], tid);
  for (int64_t i = tail0 + tid; i < kpq; i += kPostThreads) collect(keys[i], i);
  __syncthreads();
  stamp(3);  // pass 2 done
  const bool overflow = sm.n_cand > kCandCap;
  int nc = overflow ? kCandCap : sm.n_cand;
  int P = 1;
  while (P < nc) P <<= 1;
  if (P < 2) P = 2;
  const double qn2 = a.qnorm2[q];
  // ---- exact rescore: each warp takes candidates c and c + 16 together
  constexpr int kWarps = kPostThreads / 32;
  for (int c = warp; c < P; c += 2 * kWarps) {
    const int c1 = c + kWarps;
    double d0 = INFINITY, d1 = INFINITY;
    int32_t r0 = INT32_MAX, r1 = INT32_MAX;
    if (c < nc) {
      r0 = (sm.cidx[c] >> 1) * a.S + (sm.ckey[c] & 0xFF);
      const bool two = c1 < nc;
      if (two) r1 = (sm.cidx[c1] >> 1) * a.S + (sm.ckey[c1] & 0xFF);
      exact_distance2(a.sh, r0, two ? r1 : r0, two, qd, qn2, lane, d0, d1);
    }
    if (lane == 0) {
      sd_base[c] = d0;
      sr_base[c] = r0;
      if (c1 < P) {
        sd_base[c1] = d1;
        sr_base[c1] = r1;
      }
    }
  }
  __syncthreads();
  stamp(4);  // rescored
  if (P <= 32) {
    if (warp == 0) {
      double d = lane < P ? sd_base[lane] : INFINITY;
      int32_t r = lane < P ? sr_base[lane] : INT32_MAX;
      warp_sort_items(d, r, lane);
      if (lane < P) {
        sd_base[lane] = d;
        sr_base[lane] = r;
      }
    }
    __syncthreads();
  } else {
    sort_items(sd_base, sr_base, P);
  }
  const double maxd = a.max_dist ? a.max_dist[q] : INFINITY;
  int found;
  {
    int lo = 0, hi = P < a.k ? P : a.k;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (sd_base[mid] <= maxd && sd_base[mid] < INFINITY) lo = mid + 1; else hi = mid;
    }
    found = lo;
  }
  for (int i = tid; i < a.k; i += kPostThreads) {
    const bool ok = i < found;
    a.out_rows[(int64_t)q * a.k + i] = ok ? a.row_base + (int64_t)sr_base[i] * a.row_stride : -1;
    a.out_dist[(int64_t)q * a.k + i] = ok ? sd_base[i] : INFINITY;
  }
  if (tid == 0) a.out_count[q] = found;

  // ---- certificate (see finalize_kernel in k_select.cu for the argument)
  double tau;
  if (found >= a.k) tau = 1.0 - sd_base[a.k - 1];
  else if (maxd < INFINITY) tau = 1.0 - maxd;
  else tau = -INFINITY;
  const float thr = (float)(tau - (double)a.margin) - 1e-7f;
  // Every key that was not collected is < M.  If M itself still clears thr, some of those keys
  // might belong in the answer: top up with every key in [thr, M) and rescore them too.
  if (tid == 0 && overflow) sm.fail = 1;
  if (!overflow && M > kKeyNone + 1 && key_score_upper(M) >= thr) {
    const int nc0 = nc;
    auto topup = [&](int32_t kv, int64_t idx) {
      if (kv < M && kv != kKeyNone && key_score_upper(kv) >= thr) {
        const int pos = atomicAdd(&sm.n_cand, 1);
        if (pos < kCandCap) {
          sm.ckey[pos] = kv;
          sm.cidx[pos] = (int32_t)idx;
        }
      }
    };
    for (int64_t i = tid; i < n4; i += kPostThreads) {
      const int4 v = __ldg(k4 + i);
      topup(v.x, head + 4 * i);
      topup(v.y, head + 4 * i + 1);
      topup(v.z, head + 4 * i + 2);
      topup(v.w, head + 4 * i + 3);
    }
    if (tid < head) topup(keys[tid], tid);
    for (int64_t i = tail0 + tid; i < kpq; i += kPostThreads) topup(keys[i], i);
    __syncthreads();
    if (sm.n_cand > kCandCap) {
      if (tid == 0) sm.fail = 1;
    } else if (sm.n_cand > nc0) {
      // append the new rows behind the sorted candidates and sort again (rows are distinct: one key each)
      const int nc1 = sm.n_cand;
      for (int c = nc0 + warp; c < nc1; c += kPostThreads / 32) {
        const int32_t row = (sm.cidx[c] >> 1) * a.S + (sm.ckey[c] & 0xFF);
        const double d = exact_distance(a.sh, row, qd, qn2, lane);
        if (lane == 0) {
          sd_base[c] = d;
          sr_base[c] = row;
        }
      }
      int P1 = 2;
      while (P1 < nc1) P1 <<= 1;
      __syncthreads();
      for (int i = nc1 + tid; i < P1; i += kPostThreads) {
        sd_base[i] = INFINITY;
        sr_base[i] = INT32_MAX;
      }
      sort_items(sd_base, sr_base, P1);
      int lo = 0, hi = P1 < a.k ? P1 : a.k;
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (sd_base[mid] <= maxd && sd_base[mid] < INFINITY) lo = mid + 1; else hi = mid;
      }
      found = lo;
      for (int i = tid; i < a.k; i += kPostThreads) {
        const bool ok = i < found;
        a.out_rows[(int64_t)q * a.k + i] = ok ? a.row_base + (int64_t)sr_base[i] * a.row_stride : -1;
        a.out_dist[(int64_t)q * a.k + i] = ok ? sd_base[i] : INFINITY;
      }
      if (tid == 0) a.out_count[q] = found;
      nc = nc1;
    }
  }
  stamp(5);  // sorted, written, topped up
  const int64_t nsb = kpq >> 1;
  for (int64_t sb = tid; sb < nsb; sb += kPostThreads) {
    const int32_t k2 = keys[2 * sb + 1];
    if (k2 != kKeyNone && key_score_upper(k2) >= thr) {
      const int pos = atomicAdd(&sm.n_suspect, 1);
      if (pos < kSuspectCap) sm.suspects[pos] = (int32_t)sb;
    }
  }
  __syncthreads();
  stamp(6);  // certificate pass done
  const int nsus = sm.n_suspect;
  if (nsus > kSuspectCap) {
    if (tid == 0) sm.fail = 1;
  } else if (nsus > 0) {
    // ---- rescue: rescore every eligible row of each suspect sub-bucket, merge, re-emit
    const uint32_t* bm = a.bitmaps + (int64_t)(a.query_group ? a.query_group[q] : 0) * a.words_per_group;
    const int total = nsus * a.S;
    for (int t = tid; t < total; t += kPostThreads) {   // list the eligible rows first ...
      const int64_t row = (int64_t)sm.suspects[t / a.S] * a.S + (t % a.S);
      if ((bm[row >> 5] >> (row & 31)) & 1u) {
        const int pos = nc + atomicAdd(&sm.n_extra, 1);
        if (pos < kSortCap) sr_base[pos] = (int32_t)row;
      }
    }
    __syncthreads();
    const int ne = sm.n_extra;
    if (nc + ne > kSortCap) {
      if (tid == 0) sm.fail = 1;
    } else {
      for (int c = nc + warp; c < nc + ne; c += 2 * kWarps) {   // ... then rescore them two per warp
        const int c1 = c + kWarps;
        const bool two = c1 < nc + ne;
        double d0, d1;
        exact_distance2(a.sh, sr_base[c], two ? sr_base[c1] : sr_base[c], two, qd, qn2, lane, d0, d1);
        if (lane == 0) {
          sd_base[c] = d0;
          if (two) sd_base[c1] = d1;
        }
      }
      __syncthreads();
      int P2 = 2;
      while (P2 < nc + ne) P2 <<= 1;
      // the first nc entries are the sorted candidates, but sorting moved padding inside [0, P):
      // after sort_items the real entries are the first nc (padding sorts last), so this layout holds
      for (int i = nc + ne + tid; i < P2; i += kPostThreads) {
        sd_base[i] = INFINITY;
        sr_base[i] = INT32_MAX;
      }
      sort_items(sd_base, sr_base, P2);
      if (tid == 0) {
        int f2 = 0;
        int32_t prev = -1;
        for (int i = 0; i < P2 && f2 < a.k; i++) {
          if (!(sd_base[i] <= maxd) || !(sd_base[i] < INFINITY)) break;
          if (sr_base[i] == prev) continue;  // a suspect sub-bucket's keyed rows appear twice
          prev = sr_base[i];
          a.out_rows[(int64_t)q * a.k + f2] = a.row_base + (int64_t)sr_base[i] * a.row_stride;
          a.out_dist[(int64_t)q * a.k + f2] = sd_base[i];
          f2++;
        }
        a.out_count[q] = f2;
        for (int i = f2; i < a.k; i++) {
          a.out_rows[(int64_t)q * a.k + i] = -1;
          a.out_dist[(int64_t)q * a.k + i] = INFINITY;
        }
        atomicAdd(a.rescued_total, ne);
      }
    }
  }
  __syncthreads();
  stamp(7);
  if (tid == 0 && sm.fail) a.fail_list[atomicAdd(a.fail_count, 1)] = q;
}


// =====================================================================================
// Emission mode (tensor-core scorers, k <= 256): the scorer leaves a short per-query list of every
// eligible row whose approximate score cleared the query's running threshold (k_score_umma.cu).
// Nothing below T - band was ever skipped (T = the k-th best approximate score, band = 2 * margin),
// so the exact top-k is inside the list and no certificate pass over the store is needed.

constexpr int kEmitThreads = 256;
constexpr int kEmitCandMin = 512;  // entries within the band of T that one block can rescore (1024 when the query's
                                   // fp64 copy leaves room in the block's ~26 KB); more -> exact scan
constexpr int kEmitRegs = 16;   // list entries per thread held in registers during the selection

struct EmitSmem {
  unsigned long long* sk;  // [cand] (biased score << 32 | ~row): descending = best score first, smaller row first
  double* sd;              // [cand]
  int32_t* sr;             // [cand]
  int cnt[3];
  int n_keep, n_b, fail;
};

// One warp per query: the k-th largest of the seed keys (top-2 per 32 rows over the first row tiles; distinct
// rows, so k of them at or above a value bound the k-th best score of the whole store from below).
__global__ void __launch_bounds__(256) emit_seed_kernel(const int32_t* __restrict__ seed_keys, int64_t kpq,
                                                        const double* __restrict__ qnorm2, int nq, int qpad, EmitView ev) {
  const int q = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  ptx::grid_dep_launch();   // the full-store scorer may start loading; it waits for the thresholds itself
  ptx::grid_dep_wait();     // the seed keys come from the kernel before this one
  if (q >= qpad) return;
  int32_t thr = kOrderedPosInf;  // padding rows of the last query tile and queries without a cosine emit nothing
  if (q < nq) {
    const double n2 = qnorm2[q];
    if (n2 > 0.0 && n2 < INFINITY) {
      const int32_t* keys = seed_keys + (int64_t)q * kpq;
      // the keys of this query in registers (kpq <= 40 tiles x 16 = 640 = 20 per lane), one load round trip
      constexpr int kPer = 20;
      uint32_t u[kPer];
#pragma unroll
      for (int j = 0; j < kPer; j++) {
        const int64_t i = lane + 32 * j;
        u[j] = i < kpq ? ((uint32_t)(__ldg(keys + i) & ~0xFF) ^ 0x80000000u) : 0u;   // kKeyNone -> 0, below every candidate
      }
      uint32_t prefix = 0;
      for (int bit = 31; bit >= 0; bit--) {
        const uint32_t cand = prefix | (1u << bit);
        int c = 0;
#pragma unroll
        for (int j = 0; j < kPer; j++) c += u[j] >= cand;
        c = __reduce_add_sync(0xffffffffu, c);
        if (c >= ev.k) prefix = cand;
      }
      // fewer than k keys: no bound yet
      thr = prefix == 0u ? kOrderedNegInf : score_to_ordered(biased_to_score(prefix) - ev.band);
    }
  }
  if (lane == 0) {
    ev.thr[q] = thr;
    ev.count[q] = 0u;
  }
}

__global__ void __launch_bounds__(kEmitThreads, 2) emit_post_kernel(EmitPostArgs a) {
  extern __shared__ __align__(16) uint8_t raw[];
  double* qd = reinterpret_cast<double*>(raw);  // [pitch]
  __shared__ EmitSmem sm;
  const int kEmitCand = a.cand;
  if (threadIdx.x == 0) {
    sm.sk = reinterpret_cast<unsigned long long*>(qd + a.sh.pitch);
    sm.sd = reinterpret_cast<double*>(sm.sk + kEmitCand);
    sm.sr = reinterpret_cast<int32_t*>(sm.sd + kEmitCand);
  }
  __syncthreads();
  const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int kWarps = kEmitThreads / 32;
  if (a.timer && tid == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    atomicMin(a.timer, t);
  }
  const unsigned int n_raw = a.ev.count[q];
  const int n = (int)min(n_raw, (unsigned)a.ev.cap);
  unsigned long long* list = a.ev.list + (size_t)q * a.ev.cap;
  const double qn2 = a.qnorm2[q];
  const double maxd = a.max_dist ? a.max_dist[q] : INFINITY;

  for (int i = tid; i < a.sh.pitch; i += kEmitThreads) qd[i] = i < a.sh.dim ? (double)a.q[(int64_t)q * a.sh.dim + i] : 0.0;
  if (tid == 0) {
    sm.cnt[0] = sm.cnt[1] = sm.cnt[2] = 0;
    sm.n_keep = 0;
    sm.n_b = 0;
    sm.fail = n_raw > (unsigned)a.ev.cap;
  }
  // ---- the list: first 4096 entries in registers, the rest (large k only) re-read from L2
  unsigned long long e[kEmitRegs];
#pragma unroll
  for (int j = 0; j < kEmitRegs; j++) {
    const int i = tid + j * kEmitThreads;
    e[j] = i < n ? __ldcg(list + i) : 0ull;
  }
  __syncthreads();
  // ---- The query's final threshold is already (a lower bound of T) - band, so "every entry at or above it" holds
  // every row of the answer; after the scorer's last refresh that set is only a little larger than the band of T
  // itself.  Only when it does not fit the block's capacity is T found exactly (k-th largest entry, bit by bit,
  // 32 block-wide rounds) and the band cut at T - band.
  const int32_t thr_final = a.ev.thr[q];
  const uint32_t cut0 = thr_final == kOrderedPosInf ? 0xFFFFFFFFu : max(1u, (uint32_t)thr_final ^ 0x80000000u);
  {
    int c = 0;
#pragma unroll
    for (int j = 0; j < kEmitRegs; j++) c += (uint32_t)(e[j] >> 32) >= cut0;
    for (int i = kEmitRegs * kEmitThreads + tid; i < n; i += kEmitThreads) c += (uint32_t)(__ldcg(list + i) >> 32) >= cut0;
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0 && c) atomicAdd(&sm.n_b, c);
  }
  __syncthreads();
  const bool loose_fits = sm.n_b <= kEmitCand;
  __syncthreads();
  if (tid == 0) sm.n_b = 0;
  uint32_t prefix = 0;
  for (int bit = loose_fits ? -1 : 31; bit >= 0; bit--) {
    const uint32_t cand = prefix | (1u << bit);
    int c = 0;
#pragma unroll
    for (int j = 0; j < kEmitRegs; j++) c += (uint32_t)(e[j] >> 32) >= cand;
    for (int i = kEmitRegs * kEmitThreads + tid; i < n; i += kEmitThreads) c += (uint32_t)(__ldcg(list + i) >> 32) >= cand;
    c = __reduce_add_sync(0xffffffffu, c);
    int* slot = &sm.cnt[bit % 3];
    if (tid == 0) sm.cnt[(bit + 2) % 3] = 0;  // the counter of the NEXT round (bit - 1); last read two rounds ago
    if (lane == 0 && c) atomicAdd(slot, c);
    __syncthreads();
    if (*slot >= a.k) prefix = cand;
  }
  // ---- keep what lies within the band of T (everything when the list holds fewer than k rows)
  uint32_t cut = loose_fits ? cut0 : (prefix == 0u ? 1u : score_to_biased(biased_to_score(prefix) - a.ev.band));
  if (maxd < INFINITY) {  // `<= maxDistance`: a row that passes scores at least 1 - maxd - margin approximately
    const uint32_t c2 = score_to_biased((float)(1.0 - maxd) - a.margin - 1e-6f);
    cut = max(cut, c2);
  }
  auto keep = [&](unsigned long long v) {
    if ((uint32_t)(v >> 32) >= cut && v != 0ull) {
      const int pos = atomicAdd(&sm.n_keep, 1);
      if (pos < kEmitCand) sm.sk[pos] = (v & 0xFFFFFFFF00000000ull) | (unsigned long long)(~(uint32_t)v);
    }
  };
#pragma unroll
  for (int j = 0; j < kEmitRegs; j++) keep(e[j]);
  for (int i = kEmitRegs * kEmitThreads + tid; i < n; i += kEmitThreads) keep(__ldcg(list + i));
  // the list is consumed: leave it zeroed for the next batch of this lane
  for (int i = tid; i < n; i += kEmitThreads) __stcg(list + i, 0ull);
  __syncthreads();
  if (tid == 0 && a.listed_total) atomicAdd(a.listed_total, n);
  const int nk = sm.n_keep;
  if (nk > kEmitCand || sm.fail || !(qn2 > 0.0) || !(qn2 < INFINITY)) {
    const bool no_cosine = !(qn2 > 0.0) || !(qn2 < INFINITY);  // pgvector: NaN distance, never returned
    for (int i = tid; i < a.k; i += kEmitThreads) {
      a.out_rows[(int64_t)q * a.k + i] = -1;
      a.out_dist[(int64_t)q * a.k + i] = INFINITY;
    }
    if (tid == 0) {
      a.out_count[q] = 0;
      if (!no_cosine) a.fail_list[atomicAdd(a.fail_count, 1)] = q;
    }
    return;
  }
  // ---- best approximate score first
  int P = 2;
  while (P < nk) P <<= 1;
  for (int i = nk + tid; i < P; i += kEmitThreads) sm.sk[i] = 0ull;
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int i = tid; i < (P >> 1); i += kEmitThreads) {
        const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
        const bool up = ((lo & size) == 0);
        const unsigned long long x = sm.sk[lo], y = sm.sk[hi];
        if ((y > x) == up) {
          sm.sk[lo] = y;
          sm.sk[hi] = x;
        }
      }
    }
  }
  __syncthreads();
  auto row_of = [&](int i) { return (int32_t)(~(uint32_t)sm.sk[i]); };
  auto rescore = [&](int from, int to) {  // entries [from, to) of the score order -> sd / sr at the same index
    for (int c = from + warp; c < to; c += 2 * kWarps) {
      const int c1 = c + kWarps;
      const bool two = c1 < to;
      const int32_t r0 = row_of(c), r1 = two ? row_of(c1) : r0;
      double d0, d1;
      exact_distance2(a.sh, r0, r1, two, qd, qn2, lane, d0, d1);
      if (lane == 0) {
        sm.sd[c] = d0;
        sm.sr[c] = r0;
        if (two) {
          sm.sd[c1] = d1;
          sm.sr[c1] = r1;
        }
      }
    }
  };
  auto sort_first = [&](int count) {  // ascending (distance, row) over [0, count), padded to a power of two
    int Q = 2;
    while (Q < count) Q <<= 1;
    __syncthreads();
    for (int i = count + tid; i < Q; i += kEmitThreads) {
      sm.sd[i] = INFINITY;
      sm.sr[i] = INT32_MAX;
    }
    if (Q <= 32) {
      __syncthreads();
      if (warp == 0) {
        double d = lane < Q ? sm.sd[lane] : INFINITY;
        int32_t r = lane < Q ? sm.sr[lane] : INT32_MAX;
        warp_sort_items(d, r, lane);
        if (lane < Q) {
          sm.sd[lane] = d;
          sm.sr[lane] = r;
        }
      }
      __syncthreads();
    } else {
      sort_items(sm.sd, sm.sr, Q);
    }
  };
  // ---- round one: the K2 best by approximate score
  const int A = min(nk, a.K2);
  rescore(0, A);
  sort_first(A);
  // tau0 = the k-th best exact similarity so far: the final k-th is no lower, and a row that can still reach it
  // scores at least tau0 - margin approximately.  The sorted score order makes those rows a prefix of [A, nk).
  double need = A >= a.k ? 1.0 - sm.sd[a.k - 1] : -INFINITY;
  if (maxd < INFINITY) need = fmax(need, 1.0 - maxd);
  const float thr_a = (float)(need - (double)a.margin) - 1e-6f;
  {
    int c = 0;
    for (int i = A + tid; i < nk; i += kEmitThreads) c += biased_to_score((uint32_t)(sm.sk[i] >> 32)) >= thr_a;
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0 && c) atomicAdd(&sm.n_b, c);
  }
  __syncthreads();
  const int B = sm.n_b;
  if (B > 0) {
    rescore(A, A + B);  // overwrites the padding sort_first left behind [A, ...)
    sort_first(A + B);
    if (tid == 0 && a.rescued_total) atomicAdd(a.rescued_total, B);
  }
  int found;
  {
    int lo = 0, hi = min(A + B, a.k);
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (sm.sd[mid] <= maxd && sm.sd[mid] < INFINITY) lo = mid + 1; else hi = mid;
    }
    found = lo;
  }
  for (int i = tid; i < a.k; i += kEmitThreads) {
    const bool ok = i < found;
    a.out_rows[(int64_t)q * a.k + i] = ok ? a.row_base + (int64_t)sm.sr[i] * a.row_stride : -1;
    a.out_dist[(int64_t)q * a.k + i] = ok ? sm.sd[i] : INFINITY;
  }
  if (tid == 0) a.out_count[q] = found;
  if (a.timer && tid == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    atomicMax(a.timer + 1, t);
  }
}

}  // namespace

bool emit_supported(int k) { return k >= 1 && k <= 256; }

// Entries one block can hold: 1024 when they fit beside the query's fp64 copy in what a scorer CTA of the next
// batch leaves free on an SM (~26 KB), 512 otherwise.
int emit_post_capacity(int pitch) { return (size_t)pitch * 8 + 1024 * 20 + 256 <= 26 * 1024 ? 1024 : kEmitCandMin; }

int launch_emit_seed(const int32_t* seed_keys, int64_t keys_per_query, const double* qnorm2, int nq, int qpad,
                     const EmitView& ev, bool pdl, cudaStream_t st) {
  if (qpad <= 0) return 0;
  static bool attr_set = false;
  if (!attr_set) {   // the scorer's carve-out, so that its blocks can be resident beside a scorer CTA
    OMR_CUDA_OK(cudaFuncSetAttribute(emit_seed_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)((qpad + 7) / 8));
  cfg.blockDim = dim3(256);
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  OMR_CUDA_OK(cudaLaunchKernelEx(&cfg, emit_seed_kernel, seed_keys, keys_per_query, qnorm2, nq, qpad, ev));
  return 0;
}

int launch_emit_post(const EmitPostArgs& a, cudaStream_t st) {
  if (a.nq <= 0) return 0;
  if (a.K2 < a.k || a.K2 > a.cand || a.cand < kEmitCandMin)
    return fail_msg(-1, "emission post kernel: first round of %d rows for k = %d, capacity %d", a.K2, a.k, a.cand);
  const size_t smem = (size_t)a.sh.pitch * sizeof(double) + (size_t)a.cand * 20;
  static bool attr_set = false;
  if (!attr_set) {
    OMR_CUDA_OK(cudaFuncSetAttribute(emit_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 48 * 1024));
    // same carve-out as the scorer so a block fits beside a scorer CTA of the next batch (see launch_post)
    OMR_CUDA_OK(cudaFuncSetAttribute(emit_post_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    attr_set = true;
  }
  emit_post_kernel<<<a.nq, kEmitThreads, smem, st>>>(a);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

bool post_fused_supported(int K2, int k) { return K2 <= 256 && k <= 256; }

int launch_post(const PostArgs& a, cudaStream_t st) {
  if (a.nq <= 0) return 0;
  if (a.sort_cap < kCandCap || (a.sort_cap & 1)) return fail_msg(-1, "post kernel sort capacity %d is too small", a.sort_cap);
  const size_t smem = (size_t)a.sh.pitch * sizeof(double) + (size_t)a.sort_cap * 12 + sizeof(PostSmem);
  if (smem > 32 * 1024)  // static shared memory counts against the 48 KB default too
    OMR_CUDA_OK(cudaFuncSetAttribute(post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  // Same shared-memory carve-out as the scorer (maximum), so that blocks of this kernel can be resident on
  // an SM next to a scorer CTA of the following batch; an SM cannot hold two carve-out configurations.
  static bool carveout_set = false;
  if (!carveout_set) {
    OMR_CUDA_OK(cudaFuncSetAttribute(post_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    carveout_set = true;
  }
  post_kernel<<<a.nq, kPostThreads, smem, st>>>(a);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
