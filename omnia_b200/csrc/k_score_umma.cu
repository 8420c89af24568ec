// Large-batch scorer: Q * M^T on the 5th-generation tensor cores with the top-k
// pre-selection fused into the epilogue, so the Q x N score matrix never exists.
//
// Replaces, for a whole batch of queries at once, the scan behind
// `ORDER BY o.embedding <=> $q LIMIT n` (internal/memory/store.go:240-261,
// internal/memory/retrieve_multi_tier_hybrid.go:78-96) and, in its self-join
// form, the O(N^2) nested loop of BuildEntityDuplicateCandidatesQuery
// (internal/memory/consolidation/prefilter.go:112-138).
//
// Shape of one CTA (persistent, one per SM, 256 threads):
//   warp 0      TMA producer: per 64-wide K block, a 128-query x 64 box (A) and a
//               256-row x 64 box (B) land in 128B-swizzled shared memory.
//   warp 1      one lane issues tcgen05.mma (M=128 queries, N=256 rows, K=16) four
//               times per K block into one of two 256-column TMEM accumulators.
//   warp 2      allocates / frees the 512 TMEM columns; in emission mode it spends the time between on
//               the threshold refreshes: for a query whose list has grown by `trigger` entries it takes
//               the k-th best entry of the list as a new lower bound of the final k-th best score.
//   warps 4-7   epilogue: thread t owns query t of the tile (TMEM lane t), reads its
//               256 scores 32 columns at a time with tcgen05.ld, applies the scope
//               bitmap, and keeps the two best per S-row sub-bucket as packed keys.
//               In the default recall mode (k <= 256) the epilogue keeps no keys at all: it
//               compares every score with the query's running emission threshold and appends
//               the few rows that clear it to the query's candidate list (EmitEpilogue).
//   warp 3      (emission mode) drains the CTA's shared-memory ring of emissions into the per-query
//               lists in global memory and files the refresh requests, so the epilogue warps wait
//               neither for an L2 round trip nor for a refresh.
// While the epilogue drains accumulator i, the tensor core fills accumulator i^1.
//
// Rows and queries are unit-normalised before they are rounded to bf16 (k_ingest.cu),
// so the accumulator IS the approximate cosine: the norm scaling costs nothing here.
// The exact fp32->fp64 rescore of the surviving candidates is in k_select.cu.
//
// Roofline: tensor for >= ~256 queries (2*Q*N*D flops per launch), HBM below that
// (the bf16 matrix is read once per launch: consecutive CTAs take the query tiles
// of the same row tile, so its K blocks are served from L2 after the first read).
#include <algorithm>
#include <cstdlib>

#include "kernels.cuh"
#include "ptx.cuh"

namespace omr {

namespace {

constexpr int kStages = 4;
constexpr int kABytes = kQueryTile * kBlockK * 2;  // 16 KB
constexpr int kBBytes = kTileRows * kBlockK * 2;   // 32 KB
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kNumThreads = 256;
constexpr int kEpilogueWarp0 = 4;
constexpr uint32_t kTmemCols = 512;

struct __align__(8) PipeBarriers {
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

enum { kEpiTop2 = 0, kEpiPairs = 1, kEpiEmit = 2 };

constexpr int kPairStageBytes = 1024 * 8 + 512;
constexpr int kSmemBytes = kStages * kStageBytes + (int)sizeof(PipeBarriers) + kPairStageBytes + 1024;  // +1024: manual alignment

// Epilogue of the recall scorer: two best keys per sub-bucket for this thread's query.
template <int S>
struct Top2Epilogue {
  const uint32_t* bitmap_row;  // this query's filter-group bitmap, already offset to the group
  int32_t* out;                // cand_keys row of this query
  int32_t m1, m2;

  __device__ __forceinline__ void begin_tile() { m1 = kKeyNone; m2 = kKeyNone; }

  // c = 32-column chunk index inside the 256-row tile
  __device__ __forceinline__ void chunk(int row_tile, int c, const uint32_t (&v)[32]) {
    const uint32_t mask = __ldg(bitmap_row + (int64_t)row_tile * 8 + c);
    if (mask == 0xffffffffu) {
#pragma unroll
      for (int j = 0; j < 32; j++) top2_insert(m1, m2, pack_key(__uint_as_float(v[j]), (c * 32 + j) & (S - 1)));
    } else if (mask != 0u) {
#pragma unroll
      for (int j = 0; j < 32; j++) {
        const int32_t key = pack_key(__uint_as_float(v[j]), (c * 32 + j) & (S - 1));
        top2_insert(m1, m2, ((mask >> j) & 1u) ? key : kKeyNone);
      }
    }
    if (((c + 1) * 32) % S == 0) {
      const int64_t sb = (int64_t)row_tile * (kTileRows / S) + (c * 32) / S;
      *reinterpret_cast<int2*>(out + sb * 2) = make_int2(m1, m2);
      m1 = kKeyNone;
      m2 = kKeyNone;
    }
  }
};


// ---- threshold-filtered emission (recall, k <= 256) ------------------------------------
// The slot of shared memory the self-join uses to stage pairs holds, in a recall launch, the
// refresh warp's work: a ring of query numbers whose threshold is due for a refresh and the
// scratch the refresh selects in.
constexpr int kRefreshBuf = 512;
constexpr int kRefreshSlots = 64;
constexpr int kRing = 512;   // emissions staged per CTA between the epilogue warps and warp 3
struct RingEntry {
  unsigned int q1;   // query + 1, written LAST by the producer; 0 = empty
  int row;
  uint32_t bits;     // the fp32 score
};
struct RefreshQueue {
  RingEntry ring[kRing];
  uint32_t buf[kRefreshBuf];
  unsigned int slot[kRefreshSlots];  // query + 1; 0 = empty (refresh requests of the direct path)
  unsigned int tail;
  unsigned int epi_done;             // epilogue warps that have finished their tiles
  unsigned int ring_tail, ring_head;
};
static_assert(sizeof(RefreshQueue) <= kPairStageBytes, "RefreshQueue outgrew the pair-stage slot");

__device__ __forceinline__ uint32_t pick32(const uint32_t (&v)[32], int j) {
  uint32_t r = v[0];
#pragma unroll
  for (int i = 1; i < 32; i++) r = (j == i) ? v[i] : r;
  return r;
}

// Appends (score, row) to query q's list; out of line so the scorer's hot loop stays small.  Every
// `trigger`-th entry of a query (and its k-th) also files a refresh request with this CTA's warp 2.
__device__ __noinline__ void emit_entry(int32_t* thr, unsigned int* count, unsigned long long* list, int cap, int k,
                                        int trigger, RefreshQueue* rq, int q, int row, uint32_t bits) {
  const unsigned int slot = atomicAdd(count + q, 1u);
  if (slot < (unsigned)cap) {
    const unsigned long long e =
        ((unsigned long long)score_to_biased(__uint_as_float(bits)) << 32) | (unsigned long long)(uint32_t)row;
    __stcg(list + (size_t)q * cap + slot, e);
  } else if (slot == (unsigned)cap) {
    atomicMax(thr + q, kOrderedPosInf);  // list full: the exact scan answers this query, stop feeding it
  }
  const unsigned int c1 = slot + 1;
  if ((c1 & (unsigned)(trigger - 1)) == 0u || c1 == (unsigned)k) {
    const unsigned int i = atomicAdd(&rq->tail, 1u);
    atomicExch(&rq->slot[i & (kRefreshSlots - 1)], (unsigned)q + 1u);
  }
}

// What an epilogue thread does with a row that cleared its threshold: it drops (query, row, score) into the CTA's
// shared-memory ring and goes back to its tile.  Warp 3 drains the ring into the per-query lists (the global
// atomicAdd, the store, the refresh) off the epilogue's critical path: an append straight from the epilogue costs
// a round trip to L2 per entry with the whole warp waiting, which on a 125k-row shard -- where the few hundred
// entries per query all fall into a few dozen tiles -- was a third of the kernel's time.  A full ring (warp 3
// behind by 128+ entries) falls back to the direct append.
__device__ __noinline__ void emit_ring(int32_t* thr, unsigned int* count, unsigned long long* list, int cap, int k,
                                       int trigger, RefreshQueue* rq, int q, int row, uint32_t bits) {
  const unsigned int head = *reinterpret_cast<volatile unsigned int*>(&rq->ring_head);
  const unsigned int tail = *reinterpret_cast<volatile unsigned int*>(&rq->ring_tail);
  if (tail - head < (unsigned)(kRing - 128)) {   // at most 128 epilogue threads can pass this test at once
    const unsigned int i = atomicAdd(&rq->ring_tail, 1u);
    RingEntry* e = &rq->ring[i & (kRing - 1)];
    e->row = row;
    e->bits = bits;
    __threadfence_block();
    *reinterpret_cast<volatile unsigned int*>(&e->q1) = (unsigned)q + 1u;
    return;
  }
  emit_entry(thr, count, list, cap, k, trigger, rq, q, row, bits);
}

struct EmitEpilogue {
  EmitView ev;
  const uint32_t* bitmap_row;
  RefreshQueue* rq;
  int q;
  float t;  // this tile's emission threshold (the value of thr[q] when the tile began)

  __device__ __forceinline__ void begin_tile(int q_) {
    q = q_;
    t = ordered_to_score(__ldcg(ev.thr + q));
  }

  __device__ __forceinline__ void chunk(int row_tile, int c, const uint32_t (&v)[32]) {
    // a 31-step max decides first whether anything in the chunk can pass: the common case ends there, at half the
    // instructions of 32 compare-and-or steps, and no bitmap word is even loaded
    float m = __uint_as_float(v[0]);
#pragma unroll
    for (int j = 1; j < 32; j++) m = fmaxf(m, __uint_as_float(v[j]));
    if (!(m >= t)) return;
    uint32_t pass = 0;
#pragma unroll
    for (int j = 0; j < 32; j++) pass |= (__uint_as_float(v[j]) >= t) ? (1u << j) : 0u;
    pass &= __ldg(bitmap_row + (int64_t)row_tile * 8 + c);
    while (pass) {
      const int j = __ffs(pass) - 1;
      pass &= pass - 1;
      emit_ring(ev.thr, ev.count, ev.list, ev.cap, ev.k, ev.trigger, rq, q, row_tile * kTileRows + c * 32 + j, pick32(v, j));
    }
  }
};

// One refresh, whole warp: the k-th largest entry of query q's list is a lower bound of the final k-th best
// approximate score (k distinct eligible rows score at least that), so thr[q] may rise to it minus the band.
// Entries below the current cut cannot be among the k best and are skipped while the list is read.
__device__ void refresh_threshold(const EmitView& ev, int q, int lane, uint32_t* buf) {
  const unsigned int n = min(__ldcg(ev.count + q), (unsigned)ev.cap);
  if (n < (unsigned)ev.k) return;
  const int32_t cur = __ldcg(ev.thr + q);
  if (cur == kOrderedPosInf) return;
  // three quarters of the band above the cut: at least k entries lie above it (fp32 rounding of cut + band aside)
  const uint32_t filt = score_to_biased(ordered_to_score(cur) + 0.75f * ev.band);
  const unsigned long long* L = ev.list + (size_t)q * ev.cap;
  int m = 0;
  // Only the newest entries are read: they passed the highest thresholds, so they hold the best scores, and the
  // k-th largest of ANY k distinct rows is a valid bound.  Three round trips to L2 at most, however long the list.
  const unsigned int first = n > 768u ? n - 768u : 0u;
  for (unsigned int i0 = first; i0 < n && m <= kRefreshBuf - 256; i0 += 256) {
    unsigned long long e[8];   // eight loads in flight per lane: one L2 round trip per 256 entries
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const unsigned int i = i0 + u * 32 + lane;
      e[u] = i < n ? __ldcg(L + i) : 0ull;
    }
#pragma unroll
    for (int u = 0; u < 8; u++) {
      const uint32_t key = (uint32_t)(e[u] >> 32);
      const bool ok = e[u] != 0ull && key >= filt;
      const unsigned int b = __ballot_sync(0xffffffffu, ok);
      if (ok) buf[m + __popc(b & ((1u << lane) - 1u))] = key;
      m += __popc(b);
    }
  }
  __syncwarp();
  if (m < ev.k) return;
  uint32_t prefix = 0;
  for (int bit = 31; bit >= 0; bit--) {
    const uint32_t cand = prefix | (1u << bit);
    int c = 0;
    for (int j = lane; j < m; j += 32) c += buf[j] >= cand;
    c = __reduce_add_sync(0xffffffffu, c);
    if (c >= ev.k) prefix = cand;
  }
  if (lane == 0) atomicMax(ev.thr + q, score_to_ordered(biased_to_score(prefix) - ev.band));
  __syncwarp();
}

// Body of warp 3 in an emission launch: drain the ring into the per-query lists (up to four entries per lane per
// round, their atomicAdds in flight together) until the four epilogue warps are done and the ring is empty.  An
// entry that takes its query's count across a multiple of `trigger` (or to k) files a refresh request for warp 2.
// Draining and refreshing are two warps' work because at k = 100 a refresh costs as much as draining 500 entries:
// with both on one warp (ncu, 1.25M x 768, top-100) the ring stood full, two thirds of the emissions took the direct
// append with their whole epilogue warp waiting on L2, and the tensor pipe idled 43 % of the kernel.  Measured on
// one box, one warp vs two: 540 k -> 591 k queries/s there, 453 k -> 462 k at 1M x 1536 top-10, scorer 0.292 ->
// 0.281 ms on a 125k-row shard.
__device__ void drain_loop(const EmitView& ev, RefreshQueue* rq, int lane) {
  unsigned int head = 0;
  while (true) {
    unsigned int done = 0, tail = 0;
    if (lane == 0) {
      done = *reinterpret_cast<volatile unsigned int*>(&rq->epi_done);
      tail = *reinterpret_cast<volatile unsigned int*>(&rq->ring_tail);
    }
    done = __shfl_sync(0xffffffffu, done, 0);
    tail = __shfl_sync(0xffffffffu, tail, 0);
    const unsigned int avail = tail - head;
    if (!avail) {
      if (done >= 4u) break;   // ring empty: the tail was read after `done`, and the epilogue warps fence before counting
      __nanosleep(100);
      continue;
    }
    const unsigned int n = avail < 128u ? avail : 128u;
    unsigned int q1s[4], slots[4];
    int rows[4];
    uint32_t bitss[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      q1s[u] = 0u;
      const unsigned int i = (unsigned)lane + 32u * u;
      if (i < n) {
        RingEntry* e = &rq->ring[(head + i) & (kRing - 1)];
        while ((q1s[u] = *reinterpret_cast<volatile unsigned int*>(&e->q1)) == 0u) {   // index taken, entry still being written
        }
        __threadfence_block();
        rows[u] = *reinterpret_cast<volatile int*>(&e->row);
        bitss[u] = *reinterpret_cast<volatile uint32_t*>(&e->bits);
        *reinterpret_cast<volatile unsigned int*>(&e->q1) = 0u;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; u++)
      if (q1s[u]) slots[u] = atomicAdd(ev.count + (q1s[u] - 1u), 1u);
#pragma unroll
    for (int u = 0; u < 4; u++) {
      if (!q1s[u]) continue;
      const int qq = (int)q1s[u] - 1;
      if (slots[u] < (unsigned)ev.cap) {
        __stcg(ev.list + (size_t)qq * ev.cap + slots[u],
               ((unsigned long long)score_to_biased(__uint_as_float(bitss[u])) << 32) | (unsigned long long)(uint32_t)rows[u]);
      } else if (slots[u] == (unsigned)ev.cap) {
        atomicMax(ev.thr + qq, kOrderedPosInf);  // list full: the exact scan answers this query, stop feeding it
      }
      const unsigned int c1 = slots[u] + 1;
      if ((c1 & (unsigned)(ev.trigger - 1)) == 0u || c1 == (unsigned)ev.k) {
        const unsigned int i = atomicAdd(&rq->tail, 1u);
        atomicExch(&rq->slot[i & (kRefreshSlots - 1)], q1s[u]);
      }
    }
    __syncwarp();
    head += n;
    if (lane == 0) *reinterpret_cast<volatile unsigned int*>(&rq->ring_head) = head;
  }
}

// Body of warp 2 (idle once it has allocated tensor memory) in an emission launch: serve the refresh requests of
// warp 3 and of the direct path until the epilogue warps of this CTA are done -- a refresh only helps tiles that
// are still to come.  A request that is overwritten in the 64-slot queue before it is served leaves the threshold
// valid, only a little lower than it could be.
__device__ void refresh_loop(const EmitView& ev, RefreshQueue* rq, int lane) {
  while (true) {
    unsigned int done = lane == 0 ? *reinterpret_cast<volatile unsigned int*>(&rq->epi_done) : 0u;
    done = __shfl_sync(0xffffffffu, done, 0);
    if (done >= 4u) break;
    const unsigned int v0 = atomicExch(&rq->slot[lane], 0u);
    const unsigned int v1 = atomicExch(&rq->slot[lane + 32], 0u);
    // a query filed more than once in this sweep is served once (the lowest lane that holds it)
    const unsigned int same0 = __match_any_sync(0xffffffffu, v0), same1 = __match_any_sync(0xffffffffu, v1);
    unsigned int b0 = __ballot_sync(0xffffffffu, v0 != 0u && (same0 & ((1u << lane) - 1u)) == 0u);
    unsigned int b1 = __ballot_sync(0xffffffffu, v1 != 0u && (same1 & ((1u << lane) - 1u)) == 0u);
    if (!(b0 | b1)) {
      __nanosleep(200);
      continue;
    }
    // `done` is looked at before every refresh: a sweep can hold 64 requests, and the CTA must not outlive its
    // epilogue by their service time
    while (b0 | b1) {
      done = lane == 0 ? *reinterpret_cast<volatile unsigned int*>(&rq->epi_done) : 0u;
      done = __shfl_sync(0xffffffffu, done, 0);
      if (done >= 4u) return;
      unsigned int v;
      if (b0) {
        const int src = __ffs(b0) - 1;
        b0 &= b0 - 1;
        v = __shfl_sync(0xffffffffu, v0, src);
      } else {
        const int src = __ffs(b1) - 1;
        b1 &= b1 - 1;
        v = __shfl_sync(0xffffffffu, v1, src);
      }
      refresh_threshold(ev, (int)v - 1, lane, rq->buf);
    }
  }
}

// Epilogue of the self-join: emit (a, b) for a < b when the approximate cosine clears the (margin-lowered)
// floor and both rows are members of the join.
//  * A 31-step max over the 32 scores of a chunk decides first whether anything can pass; most chunks end there.
//  * Appends are aggregated per warp: one global atomicAdd per chunk that has passes, however many pass (on
//    tie-heavy corpora a chunk can pass whole), no shared-memory stage and no CTA barrier.
//  * With a LIMIT (`ORDER BY similarity DESC LIMIT n`, consolidation/prefilter.go:129-130) the join keeps a
//    running cut: passing pairs are counted in a histogram of their approximate scores (per CTA in shared
//    memory, folded into a global one by warp 3), and the lower edge T of the bin at which the count from the top
//    reaches n is a lower bound of the n-th best approximate score.  A pair of the final top n scores at least
//    T - 2 * margin approximately (same argument as the recall thresholds, kernels.cuh), so the cut
//    max(floor - margin, T - 2 * margin) never drops one, and on a corpus of near-duplicates it keeps the
//    candidate list at the size of the duplicate clusters instead of every pair above the floor.
constexpr int kJoinBins = 2048;
struct JoinHist {
  unsigned int bins[kJoinBins];
  unsigned int epi_done;
  unsigned int dirty;
};
static_assert(sizeof(JoinHist) <= kPairStageBytes, "JoinHist outgrew its shared-memory slot");

struct PairEpilogue {
  const uint32_t* member;  // bitmap over rows
  PairCand* pairs;
  unsigned long long cap;
  unsigned long long* count;
  float threshold;         // floor - margin
  // LIMIT mode (run_cut != nullptr)
  int32_t* run_cut;        // ordered-int image of the running cut, raised with atomicMax
  unsigned int* ghist;     // [kJoinBins] approximate-score histogram of the pairs emitted so far
  const uint16_t* kind;    // per row; nullptr = no `e1.kind = e2.kind` rule
  float hist_lo, hist_scale;  // bin = (score - hist_lo) * hist_scale
  float band;              // 2 * margin
  int limit;
  // per tile
  int32_t row_a;
  bool a_member;
  uint16_t kind_a;
  float cut;               // max(threshold, running cut) when the tile began
  JoinHist* hist;          // shared memory of this CTA

  __device__ __forceinline__ void begin_tile(int32_t row) {
    row_a = row;
    a_member = (__ldg(member + (row_a >> 5)) >> (row_a & 31)) & 1u;
    cut = threshold;
    if (run_cut) cut = fmaxf(cut, ordered_to_score(__ldcg(run_cut)));
    if (kind) kind_a = __ldg(kind + row_a);
  }

  __device__ __forceinline__ void chunk(int b_tile, int c, const uint32_t (&v)[32], int lane) {
    float m = __uint_as_float(v[0]);
#pragma unroll
    for (int j = 1; j < 32; j++) m = fmaxf(m, __uint_as_float(v[j]));
    const bool maybe = a_member && m >= cut;
    if (!__any_sync(0xffffffffu, maybe)) return;
    const int32_t b0 = b_tile * kTileRows + c * 32;
    uint32_t pass = 0;
    if (maybe) {
      uint32_t mask = __ldg(member + (b0 >> 5));
      // rows b <= a are the lower triangle: drop them
      if (row_a >= b0 + 31) mask = 0;
      else if (row_a >= b0) mask &= ~((2u << (row_a - b0)) - 1u);
#pragma unroll
      for (int j = 0; j < 32; j++) pass |= (__uint_as_float(v[j]) >= cut) ? (1u << j) : 0u;
      pass &= mask;
      if (kind) {
        uint32_t rest = pass;
        while (rest) {
          const int j = __ffs(rest) - 1;
          rest &= rest - 1;
          if (__ldg(kind + b0 + j) != kind_a) pass &= ~(1u << j);
        }
      }
      if (run_cut && pass) {
        uint32_t rest = pass;
        while (rest) {
          const int j = __ffs(rest) - 1;
          rest &= rest - 1;
          const int bin = (int)fminf(fmaxf((__uint_as_float(pick32(v, j)) - hist_lo) * hist_scale, 0.0f), (float)(kJoinBins - 1));
          atomicAdd(&hist->bins[bin], 1u);
        }
        hist->dirty = 1u;
      }
    }
    // warp-aggregated append
    const int mine = __popc(pass);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    if (total == 0) return;
    unsigned long long base = 0;
    if (lane == 31) base = atomicAdd(count, (unsigned long long)total);
    base = __shfl_sync(0xffffffffu, base, 31) + (unsigned long long)(incl - mine);
    while (pass) {
      const int j = __ffs(pass) - 1;
      pass &= pass - 1;
      if (base < cap) pairs[base] = PairCand{row_a, b0 + j};
      base++;
    }
  }
};

// Body of warp 3 in a LIMIT-mode join: fold this CTA's histogram into the global one and raise the running cut,
// until the four epilogue warps are done.
__device__ void join_cut_loop(const PairEpilogue& pe, JoinHist* h, int lane) {
  constexpr int kPerLane = kJoinBins / 32;
  while (true) {
    // one lane looks, every lane follows: the branch below holds warp collectives and must not diverge
    unsigned int done = 0, dirty = 0;
    if (lane == 0) {
      done = *reinterpret_cast<volatile unsigned int*>(&h->epi_done);
      dirty = *reinterpret_cast<volatile unsigned int*>(&h->dirty);
      if (dirty) h->dirty = 0u;
    }
    done = __shfl_sync(0xffffffffu, done, 0);
    dirty = __shfl_sync(0xffffffffu, dirty, 0);
    if (dirty) {
      __syncwarp();
      for (int b = lane; b < kJoinBins; b += 32) {
        const unsigned int v = atomicExch(&h->bins[b], 0u);
        if (v) atomicAdd(pe.ghist + b, v);
      }
      __threadfence();
      // lane L owns bins [kJoinBins - kPerLane*(L+1), kJoinBins - kPerLane*L): lane 0 holds the best scores
      const int hi = kJoinBins - kPerLane * lane;
      unsigned int sum = 0;
      for (int b = hi - kPerLane; b < hi; b++) sum += __ldcg(pe.ghist + b);
      unsigned int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const unsigned int who = __ballot_sync(0xffffffffu, incl >= (unsigned)pe.limit);
      if (who) {
        const int owner = __ffs(who) - 1;
        if (lane == owner) {
          unsigned int acc = incl - sum;
          int b = hi - 1;
          for (; b >= hi - kPerLane; b--) {
            acc += __ldcg(pe.ghist + b);
            if (acc >= (unsigned)pe.limit) break;
          }
          const float T = pe.hist_lo + (float)b / pe.hist_scale;   // lower edge of the bin
          atomicMax(pe.run_cut, score_to_ordered(T - pe.band));
        }
      }
      __syncwarp();
    } else {
      if (done >= 4u) break;
      __nanosleep(2000);
    }
  }
}

template <int kEpi, int S>
struct KernelArgs {
  int n_a_tiles;   // recall: query tiles; self-join: unused
  int n_b_tiles;   // 256-row tiles of the store
  int64_t n_work;  // (a tile, b tile) work items of this launch
  int64_t work_first, work_step;  // global item = work_first + w * work_step (self-join partitions; 0 / 1 otherwise)
  int sb;                         // self-join: B tiles per super-block edge (see decode_work)
  int k_blocks;
  int nq;          // real queries; tiles are padded with zero queries beyond it
#ifdef OMR_UMMA_EXPERIMENTS
  int n_stages;    // TMA ring depth in use (<= kStages)
  int exp;         // experiment bits (OMR_UMMA_EXP): 1 = epilogue only waits/arrives, 2 = MMA issue without TMA
#endif
  // recall
  const uint32_t* bitmaps;
  int words_per_group;
  const int32_t* query_group;
  int32_t* cand_keys;
  int64_t keys_per_query;
  EmitView ev;     // emission mode
  unsigned long long* timer;  // nullable: [0] earliest / [2] latest CTA start, [1] latest / [3] earliest CTA end (%globaltimer, ns)
  // self-join
  PairEpilogue pe;
};

// Debug builds (-DOMR_UMMA_EXPERIMENTS) can switch parts of the kernel off and shorten the TMA ring from
// the environment; the shipped build has neither the knobs nor the branches.
#ifdef OMR_UMMA_EXPERIMENTS
#define OMR_EXP(args, bit) ((args).exp & (bit))
#define OMR_NSTAGES(args) ((args).n_stages)
#else
#define OMR_EXP(args, bit) 0
#define OMR_NSTAGES(args) kStages
#endif

// Work item -> (A tile of 128, B tile of 256); false = nothing to do for this item.
// Recall: a-tile fastest, so the CTAs that run concurrently share B tiles through L2.
// Self-join: the tile-pair space is cut into super-blocks of (2 sb A tiles) x (sb B tiles) = sb*256 rows on
// both sides, walked column by column over the upper triangle (I <= J); inside a super-block every pair is
// an item (a-tile fastest) and the ones below the diagonal or past the end are skipped.  The two row
// blocks of a super-block (2 x sb x 786 KB bf16 at D = 1536) stay in L2 while its 2 sb^2 tile pairs are
// computed, so HBM sees each block once per super-block instead of one A tile per tile pair.
template <bool kSelfJoin>
__device__ __forceinline__ bool decode_work(int64_t w, int n_a_tiles, int n_b_tiles, int sb, int& a_tile, int& b_tile) {
  if (!kSelfJoin) {
    b_tile = (int)(w / n_a_tiles);
    a_tile = (int)(w % n_a_tiles);
    return true;
  } else {
    const int64_t per = 2ll * sb * sb;
    const int64_t blk = w / per;
    const int t = (int)(w - blk * per);
    int64_t J = (int64_t)((sqrt(8.0 * (double)blk + 1.0) - 1.0) * 0.5);
    while (J * (J + 1) / 2 > blk) J--;
    while ((J + 1) * (J + 2) / 2 <= blk) J++;
    const int64_t I = blk - J * (J + 1) / 2;
    a_tile = (int)(I * 2 * sb) + t % (2 * sb);
    b_tile = (int)(J * sb) + t / (2 * sb);
    return a_tile < n_a_tiles && b_tile < n_b_tiles && a_tile <= 2 * b_tile + 1;
  }
}

template <int kEpi, int S>
__global__ void __launch_bounds__(kNumThreads, 1)
    score_umma_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      const KernelArgs<kEpi, S> args) {
  constexpr bool kSelfJoin = kEpi == kEpiPairs;
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B tiles must start on a 1024-byte boundary.
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages * kABytes;
  PipeBarriers* bars = reinterpret_cast<PipeBarriers*>(smem + kStages * kStageBytes);
  uint8_t* aux = reinterpret_cast<uint8_t*>(bars) + ((sizeof(PipeBarriers) + 15) & ~size_t(15));
  RefreshQueue* rq = reinterpret_cast<RefreshQueue*>(aux);   // recall, emission mode
  JoinHist* jh = reinterpret_cast<JoinHist*>(aux);           // self-join, LIMIT mode

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // seed pass of the emission mode: the seed-threshold kernel and, behind it, the full-store scorer may be
  // scheduled as soon as every CTA of this kernel is resident (they wait for its results themselves)
  if (kEpi == kEpiTop2) ptx::grid_dep_launch();
  if (kSelfJoin) {
    for (int i = threadIdx.x; i < kJoinBins; i += kNumThreads) jh->bins[i] = 0u;
    if (threadIdx.x == 0) {
      jh->epi_done = 0u;
      jh->dirty = 0u;
    }
  }
  if (args.timer && threadIdx.x == 0) {
    const unsigned long long t = ptx::globaltimer();
    atomicMin(args.timer, t);
    atomicMax(args.timer + 2, t);
  }
  if (kEpi == kEpiEmit) {
    if (threadIdx.x < kRefreshSlots) rq->slot[threadIdx.x] = 0u;
    for (int i = threadIdx.x; i < kRing; i += kNumThreads) rq->ring[i].q1 = 0u;
    if (threadIdx.x == 0) {
      rq->tail = 0u;
      rq->epi_done = 0u;
      rq->ring_tail = 0u;
      rq->ring_head = 0u;
    }
  }

  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tmap_a);
    ptx::tma_prefetch_desc(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; i++) {
      ptx::mbar_init(&bars->full[i], 1);
      ptx::mbar_init(&bars->empty[i], 1);
    }
    for (int i = 0; i < 2; i++) {
      ptx::mbar_init(&bars->tmem_full[i], 1);
      ptx::mbar_init(&bars->tmem_empty[i], 128);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc(&bars->tmem_base, kTmemCols);
  ptx::tc_fence_before_sync();
  __syncthreads();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ---------------- TMA producer ----------------
    if (lane == 0 && !OMR_EXP(args, 2)) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t w = blockIdx.x; w < args.n_work; w += gridDim.x) {
        int a_tile, b_tile;
        if (!decode_work<kSelfJoin>(args.work_first + w * args.work_step, args.n_a_tiles, args.n_b_tiles, args.sb, a_tile,
                                    b_tile))
          continue;
        for (int kb = 0; kb < args.k_blocks; kb++) {
          ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&bars->full[stage], kStageBytes);
          ptx::tma_load_2d(smem_a + stage * kABytes, &tmap_a, &bars->full[stage], kb * kBlockK,
                           a_tile * kQueryTile, ptx::kEvictLast);
          ptx::tma_load_2d(smem_b + stage * kBBytes, &tmap_b, &bars->full[stage], kb * kBlockK,
                           b_tile * kTileRows, kSelfJoin ? ptx::kEvictLast : ptx::kEvictNormal);
          if (++stage == OMR_NSTAGES(args)) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer ----------------
    // Everything the MMAs take -- descriptors, accumulator address, barrier addresses -- is computed by the whole
    // warp from warp-uniform values and only the instructions themselves sit under elect.sync: computed under
    // `lane == 0`, ptxas could not prove them uniform and wrapped every tcgen05.mma in an ELECT / 5 x R2UR loop
    // (~110 instructions per K block, more cycles than the four MMAs they issue).
    constexpr uint32_t idesc = ptx::make_idesc_bf16(kQueryTile, kTileRows);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t a_base = ptx::smem_u32(smem_a), b_base = ptx::smem_u32(smem_b);
    int stage = 0;
    uint32_t phase = 0;
    uint32_t it = 0;
    for (int64_t w = blockIdx.x; w < args.n_work; w += gridDim.x) {
      if (kSelfJoin) {
        int ta, tb;
        if (!decode_work<kSelfJoin>(args.work_first + w * args.work_step, args.n_a_tiles, args.n_b_tiles, args.sb, ta, tb))
          continue;
      }
      const uint32_t as = it & 1, aphase = (it >> 1) & 1;
      it++;
      ptx::mbar_wait(&bars->tmem_empty[as], aphase ^ 1);
      ptx::tc_fence_after_sync();
      for (int kb = 0; kb < args.k_blocks; kb++) {
        if (!OMR_EXP(args, 2)) ptx::mbar_wait(&bars->full[stage], phase);
        ptx::tc_fence_after_sync();
        const uint64_t da = ptx::make_smem_desc_sw128(a_base + stage * kABytes);
        const uint64_t db = ptx::make_smem_desc_sw128(b_base + stage * kBBytes);
        const uint32_t d_tmem = tmem_u + as * kTileRows;
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < kBlockK / 16; k++) ptx::umma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          if (!OMR_EXP(args, 2)) ptx::umma_commit(&bars->empty[stage]);  // frees the smem stage when these MMAs retire
          if (kb == args.k_blocks - 1) ptx::umma_commit(&bars->tmem_full[as]);
        }
        __syncwarp();
        if (++stage == OMR_NSTAGES(args)) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (kEpi == kEpiEmit && warp == 3) {
    // ---------------- ring drain ----------------
    ptx::grid_dep_wait();   // thresholds and counts come from the seed kernel
    drain_loop(args.ev, rq, lane);
  } else if (kEpi == kEpiEmit && warp == 2) {
    // ---------------- threshold refresh (the allocator warp has nothing else to do until the end) ----------------
    ptx::grid_dep_wait();
    refresh_loop(args.ev, rq, lane);
  } else if (kSelfJoin && warp == 3) {
    if (args.pe.run_cut) join_cut_loop(args.pe, jh, lane);
  } else if (warp >= kEpilogueWarp0) {
    // ---------------- epilogue ----------------
    const int quarter = warp & 3;  // TMEM lanes [32*quarter, 32*quarter+32) are this warp's
    const int t_local = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    uint32_t it = 0;
    Top2Epilogue<S> te;
    EmitEpilogue ee;
    ee.ev = args.ev;
    ee.rq = rq;
    PairEpilogue pe = args.pe;
    pe.hist = jh;
    // the loads and the MMAs above need nothing from the seed kernel; only the thresholds read here do
    if (kEpi == kEpiEmit) ptx::grid_dep_wait();
    for (int64_t w = blockIdx.x; w < args.n_work; w += gridDim.x) {
      int a_tile, b_tile;
      if (!decode_work<kSelfJoin>(args.work_first + w * args.work_step, args.n_a_tiles, args.n_b_tiles, args.sb, a_tile,
                                  b_tile))
        continue;
      const uint32_t as = it & 1, aphase = (it >> 1) & 1;
      it++;
      if (!kSelfJoin) {
        const int q = a_tile * kQueryTile + t_local;
        const int g = (args.query_group && q < args.nq) ? __ldg(args.query_group + q) : 0;
        if (kEpi == kEpiEmit) {
          ee.bitmap_row = args.bitmaps + (int64_t)g * args.words_per_group;
          ee.begin_tile(q);
        } else {
          te.bitmap_row = args.bitmaps + (int64_t)g * args.words_per_group;
          te.out = args.cand_keys + (int64_t)q * args.keys_per_query;
          te.begin_tile();
        }
      } else {
        pe.begin_tile(a_tile * kQueryTile + t_local);
      }
      ptx::mbar_wait(&bars->tmem_full[as], aphase);
      ptx::tc_fence_after_sync();
#pragma unroll 1
      for (int c = 0; c < (OMR_EXP(args, 1) ? 0 : kTileRows / 32); c += 2) {
        uint32_t v0[32], v1[32];
        __syncwarp();  // the epilogue's data-dependent branches must reconverge before a .sync.aligned load
        ptx::tmem_ld_32x32(lane_addr + as * kTileRows + c * 32, v0);
        ptx::tmem_ld_32x32(lane_addr + as * kTileRows + (c + 1) * 32, v1);
        ptx::tmem_ld_wait();
        if (kEpi == kEpiEmit) {
          ee.chunk(b_tile, c, v0);
          ee.chunk(b_tile, c + 1, v1);
        } else if (!kSelfJoin) {
          te.chunk(b_tile, c, v0);
          te.chunk(b_tile, c + 1, v1);
        } else {
          pe.chunk(b_tile, c, v0, lane);
          pe.chunk(b_tile, c + 1, v1, lane);
        }
      }
      ptx::tc_fence_before_sync();
      ptx::mbar_arrive(&bars->tmem_empty[as]);
    }
    if (kEpi == kEpiEmit) {
      __threadfence_block();   // this warp's ring entries are visible before it counts itself done
      __syncwarp();
      if (lane == 0) atomicAdd(&rq->epi_done, 1u);
    }
    if (kSelfJoin) {
      __syncwarp();
      if (lane == 0) atomicAdd(&jh->epi_done, 1u);
    }
  }

  ptx::tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc(tmem_base, kTmemCols);
  if (args.timer && threadIdx.x == 0) {
    const unsigned long long t = ptx::globaltimer();
    atomicMax(args.timer + 1, t);
    atomicMin(args.timer + 3, t);
  }
}

// ---------------------------------------------------------------------------------------
// 2-CTA variant of the recall scorer.  A cluster of two CTAs (one TPC) works on a 256-query x
// 256-row tile: each CTA loads ITS 128 queries (A) and ITS half of the 256 rows (B), the leader
// issues one M=256 tcgen05.mma.cta_group::2 per K step, and each CTA's epilogue drains its own
// 128 TMEM lanes.  Per SM and K block this moves 32 KB through L2->smem instead of 48 KB, and
// the smaller stage allows a 6-deep TMA pipeline.
constexpr int kStages2 = 6;
constexpr int kHalfBBytes = (kTileRows / 2) * kBlockK * 2;  // 16 KB
constexpr int kStageBytes2 = kABytes + kHalfBBytes;         // 32 KB per CTA
struct __align__(8) PipeBarriers2 {
  uint64_t full[kStages2];
  uint64_t empty[kStages2];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};
constexpr int kSmemBytes2 = kStages2 * kStageBytes2 + (int)sizeof(PipeBarriers2) + kPairStageBytes + 1024;

// Work item of the pair kernel -> (A tile of 256 rows, B tile of 256 rows).  Self-join: super-blocks of
// sb x sb tiles over the upper triangle, as decode_work, at 256-row granularity on both sides.
template <bool kSelfJoin>
__device__ __forceinline__ bool decode_work2(int64_t w, int n_a_tiles, int n_b_tiles, int sb, int64_t n_clusters,
                                             int& a_tile, int& b_tile) {
  if (!kSelfJoin) {
    if (sb & 1) {
      // Row-tile-major: round r of cluster c takes row tile c + r * n_clusters and walks every query tile on it,
      // so a row tile is fetched from HBM once and re-read from L2 by the SAME cluster a few microseconds later
      // (ncu, 1M x 1536, 1024 queries: DRAM read 3.22 GB = 1.05 x the matrix, against 3.7-4.5 GB when the four
      // clusters that share a row tile are different ones; the kernel's time is the same, it is power-bound).
      // The last, partial round of row tiles is dealt out item by item so that no cluster idles.
      const int64_t i = w / n_clusters, c = w % n_clusters;
      const int64_t full = n_b_tiles / n_clusters;
      if (i < full * n_a_tiles) {
        b_tile = (int)(c + n_clusters * (i / n_a_tiles));
        a_tile = (int)(i % n_a_tiles);
        return true;
      }
      const int64_t ww = full * n_clusters * n_a_tiles + c + (i - full * n_a_tiles) * n_clusters;
      if (ww >= (int64_t)n_a_tiles * n_b_tiles) return false;
      b_tile = (int)(ww / n_a_tiles);
      a_tile = (int)(ww % n_a_tiles);
      return true;
    }
    b_tile = (int)(w / n_a_tiles);
    a_tile = (int)(w % n_a_tiles);
    return true;
  } else {
    const int64_t per = (int64_t)sb * sb;
    const int64_t blk = w / per;
    const int t = (int)(w - blk * per);
    int64_t J = (int64_t)((sqrt(8.0 * (double)blk + 1.0) - 1.0) * 0.5);
    while (J * (J + 1) / 2 > blk) J--;
    while ((J + 1) * (J + 2) / 2 <= blk) J++;
    const int64_t I = blk - J * (J + 1) / 2;
    a_tile = (int)(I * sb) + t % sb;
    b_tile = (int)(J * sb) + t / sb;
    return a_tile < n_a_tiles && b_tile < n_b_tiles && a_tile <= b_tile;
  }
}

template <int kEpi, int S>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kNumThreads, 1)
    score_umma2_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_rows,
                       const KernelArgs<kEpi, S> args) {
  constexpr bool kSelfJoin = kEpi == kEpiPairs;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + kStages2 * kABytes;
  PipeBarriers2* bars = reinterpret_cast<PipeBarriers2*>(smem + kStages2 * kStageBytes2);
  uint8_t* aux = reinterpret_cast<uint8_t*>(bars) + ((sizeof(PipeBarriers2) + 15) & ~size_t(15));
  RefreshQueue* rq = reinterpret_cast<RefreshQueue*>(aux);   // recall, emission mode
  JoinHist* jh = reinterpret_cast<JoinHist*>(aux);           // self-join, LIMIT mode

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // seed pass of the emission mode: the seed-threshold kernel and, behind it, the full-store scorer may be
  // scheduled as soon as every CTA of this kernel is resident (they wait for its results themselves)
  if (kEpi == kEpiTop2) ptx::grid_dep_launch();
  if (kSelfJoin) {
    for (int i = threadIdx.x; i < kJoinBins; i += kNumThreads) jh->bins[i] = 0u;
    if (threadIdx.x == 0) {
      jh->epi_done = 0u;
      jh->dirty = 0u;
    }
  }
  if (args.timer && threadIdx.x == 0) {
    const unsigned long long t = ptx::globaltimer();
    atomicMin(args.timer, t);
    atomicMax(args.timer + 2, t);
  }
  if (kEpi == kEpiEmit) {
    if (threadIdx.x < kRefreshSlots) rq->slot[threadIdx.x] = 0u;
    for (int i = threadIdx.x; i < kRing; i += kNumThreads) rq->ring[i].q1 = 0u;
    if (threadIdx.x == 0) {
      rq->tail = 0u;
      rq->epi_done = 0u;
      rq->ring_tail = 0u;
      rq->ring_head = 0u;
    }
  }
  const uint32_t rank = ptx::cluster_ctarank();
  const bool leader = rank == 0;
  const int64_t cluster = ptx::cluster_id_x(), n_clusters = ptx::cluster_nctaid_x();

  ptx::cluster_sync_all();  // both CTAs resident before the paired TMEM allocation
  if (warp == 0 && lane == 0) {
    ptx::tma_prefetch_desc(&tmap_q);
    ptx::tma_prefetch_desc(&tmap_rows);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages2; i++) {
      ptx::mbar_init(&bars->full[i], 2);    // both CTAs' producers arrive on the leader's barrier
      ptx::mbar_init(&bars->empty[i], 1);   // multicast commit arrives in each CTA
    }
    for (int i = 0; i < 2; i++) {
      ptx::mbar_init(&bars->tmem_full[i], 1);
      ptx::mbar_init(&bars->tmem_empty[i], 256);  // the epilogue threads of both CTAs (leader's copy is used)
    }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc_2sm(&bars->tmem_base, kTmemCols);
  ptx::tc_fence_before_sync();
  ptx::cluster_sync_all();
  ptx::tc_fence_after_sync();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ---------------- TMA producer (one per CTA) ----------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int64_t w = cluster; w < args.n_work; w += n_clusters) {
        int a_tile, b_tile;
        if (!decode_work2<kSelfJoin>(args.work_first + w * args.work_step, args.n_a_tiles, args.n_b_tiles, args.sb, n_clusters,
                                     a_tile, b_tile))
          continue;
        for (int kb = 0; kb < args.k_blocks; kb++) {
          ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(&bars->full[stage], 2 * kStageBytes2);
          else ptx::mbar_arrive_cluster(&bars->full[stage], 0);
          ptx::tma_load_2d_2sm(smem_a + stage * kABytes, &tmap_q, &bars->full[stage], kb * kBlockK,
                               a_tile * 2 * kQueryTile + (int)rank * kQueryTile, ptx::kEvictLast);
          ptx::tma_load_2d_2sm(smem_b + stage * kHalfBBytes, &tmap_rows, &bars->full[stage], kb * kBlockK,
                               b_tile * kTileRows + (int)rank * (kTileRows / 2),
                               kSelfJoin ? ptx::kEvictLast : ptx::kEvictNormal);
          if (++stage == kStages2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1 && leader) {
    // ---------------- MMA issuer (leader CTA only) ----------------
    // operands from warp-uniform values, only the instructions under elect.sync (see the 1-CTA kernel)
    constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * kQueryTile, kTileRows);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t a_base = ptx::smem_u32(smem_a), b_base = ptx::smem_u32(smem_b);
    int stage = 0;
    uint32_t phase = 0, it = 0;
    for (int64_t w = cluster; w < args.n_work; w += n_clusters) {
      if (kSelfJoin || (args.sb & 1)) {
        int ta, tb;
        if (!decode_work2<kSelfJoin>(args.work_first + w * args.work_step, args.n_a_tiles, args.n_b_tiles, args.sb, n_clusters, ta, tb))
          continue;
      }
      const uint32_t as = it & 1, aphase = (it >> 1) & 1;
      it++;
      ptx::mbar_wait(&bars->tmem_empty[as], aphase ^ 1);
      ptx::tc_fence_after_sync();
      for (int kb = 0; kb < args.k_blocks; kb++) {
        ptx::mbar_wait(&bars->full[stage], phase);
        ptx::tc_fence_after_sync();
        const uint64_t da = ptx::make_smem_desc_sw128(a_base + stage * kABytes);
        const uint64_t db = ptx::make_smem_desc_sw128(b_base + stage * kHalfBBytes);
        const uint32_t d_tmem = tmem_u + as * kTileRows;
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < kBlockK / 16; k++) ptx::umma_bf16_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
          ptx::umma_commit_2sm(&bars->empty[stage]);
          if (kb == args.k_blocks - 1) ptx::umma_commit_2sm(&bars->tmem_full[as]);
        }
        __syncwarp();
        if (++stage == kStages2) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (kEpi == kEpiEmit && warp == 3) {
    // ---------------- ring drain (one warp per CTA) ----------------
    ptx::grid_dep_wait();   // thresholds and counts come from the seed kernel
    drain_loop(args.ev, rq, lane);
  } else if (kEpi == kEpiEmit && warp == 2) {
    // ---------------- threshold refresh (the allocator warp has nothing else to do until the end) ----------------
    ptx::grid_dep_wait();
    refresh_loop(args.ev, rq, lane);
  } else if (kSelfJoin && warp == 3) {
    if (args.pe.run_cut) join_cut_loop(args.pe, jh, lane);
  } else if (warp >= kEpilogueWarp0) {
    // ---------------- epilogue (both CTAs, own TMEM lanes) ----------------
    const int quarter = warp & 3;
    const int t_local = quarter * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    uint32_t it = 0;
    Top2Epilogue<S> te;
    EmitEpilogue ee;
    ee.ev = args.ev;
    ee.rq = rq;
    PairEpilogue pe = args.pe;
    pe.hist = jh;
    // the loads and the MMAs above need nothing from the seed kernel; only the thresholds read here do
    if (kEpi == kEpiEmit) ptx::grid_dep_wait();
    for (int64_t w = cluster; w < args.n_work; w += n_clusters) {
      int a_tile, b_tile;
      if (!decode_work2<kSelfJoin>(args.work_first + w * args.work_step, args.n_a_tiles, args.n_b_tiles, args.sb, n_clusters,
                                   a_tile, b_tile))
        continue;
      const uint32_t as = it & 1, aphase = (it >> 1) & 1;
      it++;
      const int q = a_tile * 2 * kQueryTile + (int)rank * kQueryTile + t_local;
      if (!kSelfJoin) {
        const int g = (args.query_group && q < args.nq) ? __ldg(args.query_group + q) : 0;
        if (kEpi == kEpiEmit) {
          ee.bitmap_row = args.bitmaps + (int64_t)g * args.words_per_group;
          ee.begin_tile(q);
        } else {
          te.bitmap_row = args.bitmaps + (int64_t)g * args.words_per_group;
          te.out = args.cand_keys + (int64_t)q * args.keys_per_query;
          te.begin_tile();
        }
      } else {
        pe.begin_tile(q);
      }
      ptx::mbar_wait(&bars->tmem_full[as], aphase);
      ptx::tc_fence_after_sync();
#pragma unroll 1
      for (int c = 0; c < kTileRows / 32; c += 2) {
        uint32_t v0[32], v1[32];
        __syncwarp();
        ptx::tmem_ld_32x32(lane_addr + as * kTileRows + c * 32, v0);
        ptx::tmem_ld_32x32(lane_addr + as * kTileRows + (c + 1) * 32, v1);
        ptx::tmem_ld_wait();
        if (kEpi == kEpiEmit) {
          ee.chunk(b_tile, c, v0);
          ee.chunk(b_tile, c + 1, v1);
        } else if (!kSelfJoin) {
          te.chunk(b_tile, c, v0);
          te.chunk(b_tile, c + 1, v1);
        } else {
          pe.chunk(b_tile, c, v0, lane);
          pe.chunk(b_tile, c + 1, v1, lane);
        }
      }
      ptx::tc_fence_before_sync();
      ptx::mbar_arrive_cluster(&bars->tmem_empty[as], 0);
    }
    if (kEpi == kEpiEmit) {
      __threadfence_block();   // this warp's ring entries are visible before it counts itself done
      __syncwarp();
      if (lane == 0) atomicAdd(&rq->epi_done, 1u);
    }
    if (kSelfJoin) {
      __syncwarp();
      if (lane == 0) atomicAdd(&jh->epi_done, 1u);
    }
  }

  ptx::tc_fence_before_sync();
  ptx::cluster_sync_all();
  if (warp == 2) ptx::tmem_dealloc_2sm(tmem_base, kTmemCols);
  if (args.timer && threadIdx.x == 0) {
    const unsigned long long t = ptx::globaltimer();
    atomicMax(args.timer + 1, t);
    atomicMin(args.timer + 3, t);
  }
}

// Launch with the programmatic-serialization attribute when `pdl`: the kernel may start while the kernel before it
// on `st` still runs, and waits for that kernel's results itself (ptx::grid_dep_wait).
template <class Kern, class... Args>
cudaError_t launch_maybe_pdl(Kern kern, unsigned grid, unsigned block, size_t smem, cudaStream_t st, bool pdl,
                             const Args&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, args...);
}

template <int S>
int launch_recall2_s(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows_half, int nq, int n_qtiles256, int n_tiles,
                     int k_blocks, const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                     int32_t* cand_keys, int num_sms, cudaStream_t st) {
  KernelArgs<kEpiTop2, S> a{};
  a.n_a_tiles = n_qtiles256;
  a.n_b_tiles = n_tiles;
  a.n_work = (int64_t)n_qtiles256 * n_tiles;
  a.work_first = 0;
  a.work_step = 1;
  a.k_blocks = k_blocks;
  a.nq = nq;
  a.bitmaps = bitmaps;
  a.words_per_group = words_per_group;
  a.query_group = query_group;
  a.cand_keys = cand_keys;
  a.keys_per_query = (int64_t)n_tiles * (kTileRows / S) * 2;
  a.sb = 1;      // row-tile-major order (decode_work2)
  auto kern = score_umma2_kernel<kEpiTop2, S>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes2));
  int64_t clusters = num_sms / 2;
  if (clusters > a.n_work) clusters = a.n_work;
  kern<<<(unsigned)(2 * clusters), kNumThreads, kSmemBytes2, st>>>(tmap_q, tmap_rows_half, a);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

template <int S>
int launch_recall_s(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows, int nq, int n_qtiles, int n_tiles, int k_blocks,
                    const uint32_t* bitmaps, int words_per_group, const int32_t* query_group, int32_t* cand_keys,
                    int num_sms, cudaStream_t st) {
  KernelArgs<kEpiTop2, S> a{};
  a.n_a_tiles = n_qtiles;
  a.n_b_tiles = n_tiles;
  a.n_work = (int64_t)n_qtiles * n_tiles;
  a.work_first = 0;
  a.work_step = 1;
  a.k_blocks = k_blocks;
  a.nq = nq;
  a.bitmaps = bitmaps;
  a.words_per_group = words_per_group;
  a.query_group = query_group;
  a.cand_keys = cand_keys;
  a.keys_per_query = (int64_t)n_tiles * (kTileRows / S) * 2;
#ifdef OMR_UMMA_EXPERIMENTS
  a.n_stages = kStages;
  a.exp = 0;
  if (const char* e = getenv("OMR_UMMA_EXP")) a.exp = atoi(e);
  if (const char* e = getenv("OMR_UMMA_STAGES")) a.n_stages = std::max(2, std::min(kStages, atoi(e)));
#endif
  auto kern = score_umma_kernel<kEpiTop2, S>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int grid = num_sms;
  if ((int64_t)grid > a.n_work) grid = (int)a.n_work;
  kern<<<grid, kNumThreads, kSmemBytes, st>>>(tmap_q, tmap_rows, a);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

int launch_score_umma_emit(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows, int nq, int n_qtiles, int n_tiles,
                           int k_blocks, const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                           const EmitView& ev, unsigned long long* timer, int num_sms, bool pdl, cudaStream_t st) {
  if (n_qtiles <= 0 || n_tiles <= 0) return 0;
  KernelArgs<kEpiEmit, 256> a{};
  a.n_a_tiles = n_qtiles;
  a.n_b_tiles = n_tiles;
  a.n_work = (int64_t)n_qtiles * n_tiles;
  a.work_first = 0;
  a.work_step = 1;
  a.k_blocks = k_blocks;
  a.nq = nq;
  a.bitmaps = bitmaps;
  a.words_per_group = words_per_group;
  a.query_group = query_group;
  a.ev = ev;
  a.timer = timer;
#ifdef OMR_UMMA_EXPERIMENTS
  a.n_stages = kStages;
  a.exp = 0;
  if (const char* e = getenv("OMR_UMMA_EXP")) a.exp = atoi(e);
#endif
  auto kern = score_umma_kernel<kEpiEmit, 256>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int grid = num_sms;
  if ((int64_t)grid > a.n_work) grid = (int)a.n_work;
  OMR_CUDA_OK(launch_maybe_pdl(kern, (unsigned)grid, kNumThreads, kSmemBytes, st, pdl, tmap_q, tmap_rows, a));
  return 0;
}

int launch_score_umma2_emit(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows_half, int nq, int n_qtiles256,
                            int n_tiles, int k_blocks, const uint32_t* bitmaps, int words_per_group,
                            const int32_t* query_group, const EmitView& ev, unsigned long long* timer, int num_sms,
                            bool pdl, cudaStream_t st) {
  if (n_qtiles256 <= 0 || n_tiles <= 0) return 0;
  KernelArgs<kEpiEmit, 256> a{};
  a.n_a_tiles = n_qtiles256;
  a.n_b_tiles = n_tiles;
  a.n_work = (int64_t)n_qtiles256 * n_tiles;
  a.work_first = 0;
  a.work_step = 1;
  a.k_blocks = k_blocks;
  a.nq = nq;
  a.bitmaps = bitmaps;
  a.words_per_group = words_per_group;
  a.query_group = query_group;
  a.ev = ev;
  a.sb = 1;      // row-tile-major order (decode_work2)
  a.timer = timer;
  auto kern = score_umma2_kernel<kEpiEmit, 256>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes2));
  int64_t clusters = num_sms / 2;
  if (clusters > a.n_work) clusters = a.n_work;
  OMR_CUDA_OK(launch_maybe_pdl(kern, (unsigned)(2 * clusters), kNumThreads, kSmemBytes2, st, pdl, tmap_q, tmap_rows_half, a));
  return 0;
}

int score_umma_smem_bytes() { return kSmemBytes; }

int launch_score_umma(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows, int nq, int n_qtiles, int n_tiles,
                      int k_blocks, int S, const uint32_t* bitmaps, int words_per_group,
                      const int32_t* query_group, int32_t* cand_keys, int num_sms, cudaStream_t st) {
  if (n_qtiles <= 0 || n_tiles <= 0) return 0;
  switch (S) {
    case 32: return launch_recall_s<32>(tmap_q, tmap_rows, nq, n_qtiles, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    case 64: return launch_recall_s<64>(tmap_q, tmap_rows, nq, n_qtiles, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    case 128: return launch_recall_s<128>(tmap_q, tmap_rows, nq, n_qtiles, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    case 256: return launch_recall_s<256>(tmap_q, tmap_rows, nq, n_qtiles, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  }
  return fail_msg(-2, "sub-bucket size must be 32, 64, 128 or 256, got %d", S);
}

// tmap_rows_half: the row matrix with a 128-row box (each CTA of a pair loads half a row tile).
int launch_score_umma2(const CUtensorMap& tmap_q, const CUtensorMap& tmap_rows_half, int nq, int n_qtiles256, int n_tiles,
                       int k_blocks, int S, const uint32_t* bitmaps, int words_per_group,
                       const int32_t* query_group, int32_t* cand_keys, int num_sms, cudaStream_t st) {
  if (n_qtiles256 <= 0 || n_tiles <= 0) return 0;
  switch (S) {
    case 32: return launch_recall2_s<32>(tmap_q, tmap_rows_half, nq, n_qtiles256, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    case 64: return launch_recall2_s<64>(tmap_q, tmap_rows_half, nq, n_qtiles256, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    case 128: return launch_recall2_s<128>(tmap_q, tmap_rows_half, nq, n_qtiles256, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    case 256: return launch_recall2_s<256>(tmap_q, tmap_rows_half, nq, n_qtiles256, n_tiles, k_blocks, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  }
  return fail_msg(-2, "sub-bucket size must be 32, 64, 128 or 256, got %d", S);
}

// Work items of one partition of the join's tile-pair space (super-blocks over the upper triangle, see
// decode_work / decode_work2); a launch may cover any contiguous range of them.
static int join_sb(int k_blocks) {
  // super-block edge: two row blocks of sb * 256 rows must sit in L2 together (126 MB, two dies); 24 tiles =
  // 18.9 MB per block at pitch 1536, scaled up for narrower rows
  int sb = 24;
  if (k_blocks > 0) sb = std::max(8, std::min(96, 24 * 24 / k_blocks));
  if (const char* e = getenv("OMR_SELFJOIN_SB")) sb = std::max(1, atoi(e));
  return sb;
}

int64_t selfjoin_items(int64_t n, int k_blocks, int part_index, int part_count, bool pair_kernel) {
  if (n <= 0) return 0;
  const int64_t nb = (n + kTileRows - 1) / kTileRows;
  const int sb = join_sb(k_blocks);
  const int64_t NB = (nb + sb - 1) / sb;
  const int64_t total = NB * (NB + 1) / 2 * ((pair_kernel ? 1ll : 2ll) * sb * sb);
  if (part_count < 1) part_count = 1;
  return part_index < total ? (total - part_index + part_count - 1) / part_count : 0;
}

template <typename Args>
static void fill_join_args(Args& a, const JoinLaunch& j, int64_t items) {
  const int part_count = j.part_count < 1 ? 1 : j.part_count;
  const int64_t first = std::max<int64_t>(0, j.chunk_first);
  const int64_t count = j.chunk_items < 0 ? items - first : std::min(j.chunk_items, items - first);
  a.sb = join_sb(j.k_blocks);
  a.work_first = j.part_index + first * part_count;   // partition p of P takes the items p, p+P, p+2P, ... (interleaved:
  a.work_step = part_count;                           // balanced); a chunk is a contiguous range of those
  a.n_work = std::max<int64_t>(0, count);
  a.k_blocks = j.k_blocks;
  a.pe.member = j.member_bitmap;
  a.pe.pairs = j.pairs;
  a.pe.cap = j.cap;
  a.pe.count = j.pair_count;
  a.pe.threshold = j.threshold;
  a.pe.run_cut = j.limit > 0 ? j.run_cut : nullptr;
  a.pe.ghist = j.ghist;
  a.pe.kind = j.kind;
  a.pe.hist_lo = j.hist_lo;
  a.pe.hist_scale = j.hist_scale;
  a.pe.band = j.band;
  a.pe.limit = j.limit;
}

// Self-join on CTA pairs: tmap_half is the 128-row-box map of the member matrix (it serves both operands:
// each CTA loads its half of the 256-row A tile and its half of the 256-row B tile).
int launch_selfjoin_umma2(const CUtensorMap& tmap_half, const JoinLaunch& j, int num_sms, cudaStream_t st) {
  if (j.n <= 0) return 0;
  const int64_t nb = (j.n + kTileRows - 1) / kTileRows;
  KernelArgs<kEpiPairs, 256> a{};
  a.n_a_tiles = (int)nb;
  a.n_b_tiles = (int)nb;
  fill_join_args(a, j, selfjoin_items(j.n, j.k_blocks, j.part_index, j.part_count, true));
  if (a.n_work == 0) return 0;
  auto kern = score_umma2_kernel<kEpiPairs, 256>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes2));
  int64_t clusters = num_sms / 2;
  if (clusters > a.n_work) clusters = a.n_work;
  kern<<<(unsigned)(2 * clusters), kNumThreads, kSmemBytes2, st>>>(tmap_half, tmap_half, a);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_selfjoin_umma(const CUtensorMap& tmap_a, const CUtensorMap& tmap_b, const JoinLaunch& j, int num_sms,
                         cudaStream_t st) {
  if (j.n <= 0) return 0;
  const int64_t nb = (j.n + kTileRows - 1) / kTileRows;
  KernelArgs<kEpiPairs, 256> a{};
  a.n_a_tiles = (int)(2 * nb);
  a.n_b_tiles = (int)nb;
  fill_join_args(a, j, selfjoin_items(j.n, j.k_blocks, j.part_index, j.part_count, false));
  if (a.n_work == 0) return 0;
#ifdef OMR_UMMA_EXPERIMENTS
  a.n_stages = kStages;
  a.exp = 0;
#endif
  auto kern = score_umma_kernel<kEpiPairs, 256>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
  int grid = num_sms;
  if ((int64_t)grid > a.n_work) grid = (int)a.n_work;
  kern<<<grid, kNumThreads, kSmemBytes, st>>>(tmap_a, tmap_b, a);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
