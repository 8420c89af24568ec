// Small-batch scorer: an HBM-streaming "GEMV" for 1..8 queries.
//
// Replaces the per-row cosine_distance calls a pgvector sequential scan makes for
// `ORDER BY o.embedding <=> $1 LIMIT k` (internal/memory/store_read.go:266-280,
// internal/memory/store.go:240-261) when only a handful of queries are in flight
// (dedup-on-write, batch-1 recall, BASELINE config C5).
//
// Roofline: HBM.  Algorithmic bytes per launch = n_tiles * 256 * pitch * 2 (the
// bf16 matrix is read exactly once, 16 bytes per lane per load, fully coalesced)
// + the scope bitmap.  Queries live in shared memory for the whole kernel; the
// only global writes are two packed keys per (query, sub-bucket).
#include "kernels.cuh"
#include "ptx.cuh"

namespace omr {

namespace {

constexpr int kRowsPerIter = 4;

__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__device__ __forceinline__ float bf16_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

__device__ __forceinline__ float dot8(const uint4& x, const float4& a, const float4& b, float acc) {
  acc = fmaf(bf16_lo(x.x), a.x, acc);
  acc = fmaf(bf16_hi(x.x), a.y, acc);
  acc = fmaf(bf16_lo(x.y), a.z, acc);
  acc = fmaf(bf16_hi(x.y), a.w, acc);
  acc = fmaf(bf16_lo(x.z), b.x, acc);
  acc = fmaf(bf16_hi(x.z), b.y, acc);
  acc = fmaf(bf16_lo(x.w), b.z, acc);
  acc = fmaf(bf16_hi(x.w), b.w, acc);
  return acc;
}

// 256 threads = 8 warps; warp w scores rows [32w, 32w+32) of each 256-row tile,
// kRowsPerIter rows at a time so every query fragment read from shared memory is
// reused across that many rows.  Shared layout of a query: element
// e = (c*32 + lane)*8 + h*4 + j sits at float4 index ((c*2 + h)*32 + lane), so a
// warp's LDS.128 is one conflict-free 512-byte line.
template <int NQ>
__global__ void __launch_bounds__(256, 2)
    score_stream_kernel(const __nv_bfloat16* __restrict__ rows, int pitch, int n_tiles,
                        const __nv_bfloat16* __restrict__ qn, int nq, int S,
                        const uint32_t* __restrict__ bitmaps, int words_per_group,
                        const int32_t* __restrict__ query_group, int32_t* __restrict__ cand_keys,
                        int64_t keys_per_query) {
  extern __shared__ __align__(16) float smem[];
  const int chunks = (pitch + 255) / 256;
  float* qs = smem;                       // [NQ][chunks][2][32][4]
  float* sc = qs + NQ * chunks * 256;     // [NQ][256]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  for (int idx = tid; idx < NQ * chunks * 256; idx += 256) {
    const int qi = idx / (chunks * 256), e = idx % (chunks * 256);
    float v = 0.0f;
    if (qi < nq && e < pitch) v = __bfloat162float(qn[(int64_t)qi * pitch + e]);
    const int c = e >> 8, rem = e & 255;
    qs[((qi * chunks + c) * 2 + ((rem & 7) >> 2)) * 128 + (rem >> 3) * 4 + (rem & 3)] = v;
  }
  __syncthreads();
  const float4* qs4 = reinterpret_cast<const float4*>(qs);

  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int64_t base = (int64_t)tile * kTileRows + warp * 32;
#pragma unroll 1
    for (int rr = 0; rr < 32; rr += kRowsPerIter) {
      float acc[kRowsPerIter][NQ];
#pragma unroll
      for (int r = 0; r < kRowsPerIter; r++)
#pragma unroll
        for (int qi = 0; qi < NQ; qi++) acc[r][qi] = 0.0f;
      const __nv_bfloat16* rp = rows + (base + rr) * pitch;
#pragma unroll 2
      for (int c = 0; c < chunks; c++) {
        const int e0 = (c * 32 + lane) * 8;
        if (e0 < pitch) {
          uint4 x[kRowsPerIter];
#pragma unroll
          for (int r = 0; r < kRowsPerIter; r++) x[r] = ldg_stream(rp + (int64_t)r * pitch + e0);
#pragma unroll
          for (int qi = 0; qi < NQ; qi++) {
            const float4 a = qs4[((qi * chunks + c) * 2 + 0) * 32 + lane];
            const float4 b = qs4[((qi * chunks + c) * 2 + 1) * 32 + lane];
#pragma unroll
            for (int r = 0; r < kRowsPerIter; r++) acc[r][qi] = dot8(x[r], a, b, acc[r][qi]);
          }
        }
      }
#pragma unroll
      for (int r = 0; r < kRowsPerIter; r++)
#pragma unroll
        for (int qi = 0; qi < NQ; qi++) {
          float v = acc[r][qi];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          if (lane == 0) sc[qi * 256 + warp * 32 + rr + r] = v;
        }
    }
    __syncthreads();
    // Warp qi reduces query qi's 256 scores to two keys per S-row sub-bucket.
    if (warp < NQ && warp < nq) {
      const int qi = warp;
      const int g = query_group ? query_group[qi] : 0;
      const uint32_t word = bitmaps[(int64_t)g * words_per_group + tile * 8 + (lane >> 2)];
      const uint32_t bits = (word >> ((lane & 3) * 8)) & 0xFFu;
      int32_t m1 = kKeyNone, m2 = kKeyNone;
#pragma unroll
      for (int j = 0; j < 8; j++) {
        const int rt = lane * 8 + j;
        const int32_t key = ((bits >> j) & 1u) ? pack_key(sc[qi * 256 + rt], rt & (S - 1)) : kKeyNone;
        top2_insert(m1, m2, key);
      }
      const int lanes_per_sb = S >> 3;
      for (int o = 1; o < lanes_per_sb; o <<= 1) {
        const int32_t o1 = __shfl_xor_sync(0xffffffffu, m1, o);
        const int32_t o2 = __shfl_xor_sync(0xffffffffu, m2, o);
        top2_insert(m1, m2, o1);
        top2_insert(m1, m2, o2);
      }
      if ((lane & (lanes_per_sb - 1)) == 0) {
        const int sb = tile * (kTileRows / S) + lane / lanes_per_sb;
        int32_t* dst = cand_keys + (int64_t)qi * keys_per_query + (int64_t)sb * 2;
        dst[0] = m1;
        dst[1] = m2;
      }
    }
    __syncthreads();
  }
}


// ---------------------------------------------------------------------------------------
// Single-query scorer with bulk-copy staging.  One producer thread streams the matrix through a
// ring of shared-memory stages with cp.async.bulk (a tile's rows are contiguous in memory, so no
// tensor map is needed); eight consumer warps each take one row of a stage at a time, with the
// query held in registers.  The copy engine keeps ~150 KB in flight per SM, which is what it
// takes to reach HBM speed; the LDG-based kernel above tops out lower because its loads are
// bounded by registers.
constexpr int kBulkMaxStages = 6;
constexpr int kBulkConsumerWarps = 8;
constexpr int kBulkThreads = (kBulkConsumerWarps + 1) * 32;

struct __align__(8) BulkBarriers {
  uint64_t full[kBulkMaxStages];
  uint64_t empty[kBulkMaxStages];
};

// CHUNKS = 16-byte chunks per lane per row = ceil(pitch / 256); R = rows a warp scores together
// (a stage holds 8*R rows), which gives the FMA and shuffle chains R-way instruction parallelism.
template <int CHUNKS, int R>
__global__ void __launch_bounds__(kBulkThreads, 1)
    score_stream1_kernel(const __nv_bfloat16* __restrict__ rows, int pitch, int n_tiles, int n_stages,
                         const __nv_bfloat16* __restrict__ qn, int S, const uint32_t* __restrict__ bitmaps,
                         int words_per_group, const int32_t* __restrict__ query_group,
                         int32_t* __restrict__ cand_keys) {
  extern __shared__ __align__(128) uint8_t bsm[];
  constexpr int kRowsPerStage = kBulkConsumerWarps * R;
  const int stage_bytes = kRowsPerStage * pitch * 2;
  BulkBarriers* bars = reinterpret_cast<BulkBarriers*>(bsm + (size_t)n_stages * stage_bytes);
  int32_t* part = reinterpret_cast<int32_t*>(bars + 1);  // [kBulkConsumerWarps][2] partial top-2 of a sub-bucket
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int i = 0; i < n_stages; i++) {
      ptx::mbar_init(&bars->full[i], 1);
      ptx::mbar_init(&bars->empty[i], kBulkConsumerWarps);
    }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  constexpr int kStagesPerTile = kTileRows / kRowsPerStage;

  if (warp == kBulkConsumerWarps) {
    // ---------------- producer ----------------
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const uint8_t* src = reinterpret_cast<const uint8_t*>(rows + (int64_t)tile * kTileRows * pitch);
        for (int st = 0; st < kStagesPerTile; st++) {
          ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
          ptx::mbar_arrive_expect_tx(&bars->full[stage], stage_bytes);
          ptx::bulk_load_1d(bsm + (size_t)stage * stage_bytes, src + (size_t)st * stage_bytes, stage_bytes,
                            &bars->full[stage], ptx::kEvictFirst);
          if (++stage == n_stages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
    return;
  }

  // ---------------- consumers ----------------
  const uint32_t* bitmap = bitmaps + (int64_t)(query_group ? __ldg(query_group) : 0) * words_per_group;
  // query fragment of this lane: elements (c*32 + lane)*8 .. +8 for each chunk c
  float qf[CHUNKS][8];
#pragma unroll
  for (int c = 0; c < CHUNKS; c++) {
    const int e0 = (c * 32 + lane) * 8;
#pragma unroll
    for (int j = 0; j < 8; j++) qf[c][j] = e0 + j < pitch ? __bfloat162float(qn[e0 + j]) : 0.0f;
  }
  int stage = 0;
  uint32_t phase = 0;
  int32_t m1 = kKeyNone, m2 = kKeyNone;
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    for (int st = 0; st < kStagesPerTile; st++) {
      ptx::mbar_wait(&bars->full[stage], phase);
      const uint8_t* sbase = bsm + (size_t)stage * stage_bytes;
      // warp w scores rows w, w+8, ..., w+8(R-1) of the stage together
      float acc[R];
#pragma unroll
      for (int r = 0; r < R; r++) acc[r] = 0.0f;
#pragma unroll
      for (int c = 0; c < CHUNKS; c++) {
        const int i = c * 32 + lane;
        if (i * 8 < pitch) {
          uint4 x[R];
#pragma unroll
          for (int r = 0; r < R; r++)
            x[r] = reinterpret_cast<const uint4*>(sbase + (size_t)(warp + kBulkConsumerWarps * r) * pitch * 2)[i];
#pragma unroll
          for (int r = 0; r < R; r++) {
            acc[r] = fmaf(bf16_lo(x[r].x), qf[c][0], acc[r]);
            acc[r] = fmaf(bf16_hi(x[r].x), qf[c][1], acc[r]);
            acc[r] = fmaf(bf16_lo(x[r].y), qf[c][2], acc[r]);
            acc[r] = fmaf(bf16_hi(x[r].y), qf[c][3], acc[r]);
            acc[r] = fmaf(bf16_lo(x[r].z), qf[c][4], acc[r]);
            acc[r] = fmaf(bf16_hi(x[r].z), qf[c][5], acc[r]);
            acc[r] = fmaf(bf16_lo(x[r].w), qf[c][6], acc[r]);
            acc[r] = fmaf(bf16_hi(x[r].w), qf[c][7], acc[r]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bars->empty[stage]);   // the stage's bytes are in registers now
      if (++stage == n_stages) {
        stage = 0;
        phase ^= 1;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int r = 0; r < R; r++) acc[r] += __shfl_xor_sync(0xffffffffu, acc[r], o);
      }
      const int rt0 = st * kRowsPerStage;  // kRowsPerStage <= 32 rows: one bitmap word
      const uint32_t word = __ldg(bitmap + (int64_t)tile * 8 + (rt0 >> 5)) >> (rt0 & 31);
#pragma unroll
      for (int r = 0; r < R; r++) {
        const int rs = warp + kBulkConsumerWarps * r;
        if ((word >> rs) & 1u) top2_insert(m1, m2, pack_key(acc[r], (rt0 + rs) & (S - 1)));
      }
      // sub-bucket boundary: merge the eight warps' partial top-2 and emit two keys
      if (((st + 1) * kRowsPerStage) % S == 0) {
        if (lane == 0) {
          part[warp * 2] = m1;
          part[warp * 2 + 1] = m2;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kBulkConsumerWarps * 32) : "memory");
        if (warp == 0) {
          int32_t a = lane < kBulkConsumerWarps * 2 ? part[lane] : kKeyNone, b = kKeyNone;
#pragma unroll
          for (int o = 1; o < 16; o <<= 1) {
            const int32_t oa = __shfl_xor_sync(0xffffffffu, a, o), ob = __shfl_xor_sync(0xffffffffu, b, o);
            top2_insert(a, b, oa);
            top2_insert(a, b, ob);
          }
          if (lane == 0) {
            const int64_t sb = (int64_t)tile * (kTileRows / S) + ((st + 1) * kRowsPerStage) / S - 1;
            *reinterpret_cast<int2*>(cand_keys + sb * 2) = make_int2(a, b);
          }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kBulkConsumerWarps * 32) : "memory");
        m1 = kKeyNone;
        m2 = kKeyNone;
      }
    }
  }
}

template <int CHUNKS, int R>
int launch_stream1_r(const ShardView& sh, const __nv_bfloat16* qn, int n_tiles, int S, const uint32_t* bitmaps,
                     int words_per_group, const int32_t* query_group, int32_t* cand_keys, int num_sms,
                     cudaStream_t st) {
  const size_t stage_bytes = (size_t)kBulkConsumerWarps * R * sh.pitch * 2;
  int n_stages = (int)((200 * 1024) / stage_bytes);
  if (n_stages > kBulkMaxStages) n_stages = kBulkMaxStages;
  const size_t smem = n_stages * stage_bytes + sizeof(BulkBarriers) + kBulkConsumerWarps * 2 * 4;
  auto kern = score_stream1_kernel<CHUNKS, R>;
  OMR_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = num_sms;
  if (grid > n_tiles) grid = n_tiles;
  kern<<<grid, kBulkThreads, smem, st>>>(sh.rows_bf16, sh.pitch, n_tiles, n_stages, qn, S, bitmaps, words_per_group,
                                         query_group, cand_keys);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

// 32 rows per stage (4 per warp) when two such stages fit in shared memory, else 16.
template <int CHUNKS>
int launch_stream1(const ShardView& sh, const __nv_bfloat16* qn, int n_tiles, int S, const uint32_t* bitmaps,
                   int words_per_group, const int32_t* query_group, int32_t* cand_keys, int num_sms,
                   cudaStream_t st) {
  if ((size_t)2 * 32 * sh.pitch * 2 <= 200 * 1024)
    return launch_stream1_r<CHUNKS, 4>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  return launch_stream1_r<CHUNKS, 2>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
}

template <int NQ>
int launch_nq(const ShardView& sh, const __nv_bfloat16* qn, int nq, int n_tiles, int S, const uint32_t* bitmaps,
              int words_per_group, const int32_t* query_group, int32_t* cand_keys, int num_sms, cudaStream_t st) {
  const int chunks = (sh.pitch + 255) / 256;
  const size_t smem = (size_t)(NQ * chunks * 256 + NQ * 256) * sizeof(float);
  OMR_CUDA_OK(cudaFuncSetAttribute(score_stream_kernel<NQ>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int grid = num_sms * 2;
  if (grid > n_tiles) grid = n_tiles;
  const int64_t kpq = (int64_t)n_tiles * (kTileRows / S) * 2;
  score_stream_kernel<NQ><<<grid, 256, smem, st>>>(sh.rows_bf16, sh.pitch, n_tiles, qn, nq, S, bitmaps,
                                                    words_per_group, query_group, cand_keys, kpq);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace

bool g_force_ldg_stream = false;  // OMR_STREAM_LDG=1 keeps the LDG-based kernel for a single query (A/B measurements)

int launch_score_stream(const ShardView& sh, const __nv_bfloat16* qn, int nq, int n_tiles, int S,
                        const uint32_t* bitmaps, int words_per_group, const int32_t* query_group,
                        int32_t* cand_keys, int num_sms, cudaStream_t st) {
  if (n_tiles <= 0) return 0;
  if (nq <= 1 && !g_force_ldg_stream) {
    switch ((sh.pitch + 255) / 256) {
      case 1: return launch_stream1<1>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      case 2: return launch_stream1<2>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      case 3: return launch_stream1<3>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      case 4: return launch_stream1<4>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      case 5: return launch_stream1<5>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      case 6: return launch_stream1<6>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      case 7: return launch_stream1<7>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
      default: return launch_stream1<8>(sh, qn, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
    }
  }
  if (nq <= 1) return launch_nq<1>(sh, qn, nq, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  if (nq <= 2) return launch_nq<2>(sh, qn, nq, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  if (nq <= 4) return launch_nq<4>(sh, qn, nq, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  if (nq <= 8) return launch_nq<8>(sh, qn, nq, n_tiles, S, bitmaps, words_per_group, query_group, cand_keys, num_sms, st);
  return fail_msg(-2, "score_stream handles at most 8 queries, got %d", nq);
}

}  // namespace omr
