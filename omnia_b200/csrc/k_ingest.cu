// Ingest-side kernels: write a batch of fp32 vectors into the resident matrices,
// prepare a query batch, and evaluate the recall WHERE clause into bitmaps.
//
// Reference rows: a9 (UpdateEmbedding is the ingest edge,
// internal/memory/store_write.go:557-600) and a8 (the eligibility predicate,
// internal/memory/store.go:250-258, store_read.go:271-278,
// retrieve_multi_tier.go:313-328).
#include "kernels.cuh"

namespace omr {

namespace {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One warp per vector.  Writes the fp32 master row (zero padded to `pitch`), the
// fp64 squared norm, and the unit-normalised bf16 row the scorers stream.
// HBM-bound: 4*dim bytes in, 6*pitch bytes out per row.
__global__ void __launch_bounds__(256) normalize_rows_kernel(const float* __restrict__ src, int64_t n, int dim,
                                                             int pitch, int64_t row0,
                                                             const int64_t* __restrict__ dst_rows,
                                                             __nv_bfloat16* __restrict__ out_bf16,
                                                             float* __restrict__ out_f32,
                                                             double* __restrict__ norm2, int64_t n_pad,
                                                             int32_t* __restrict__ bad_flag) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= n_pad) return;
  const int64_t dst = (dst_rows && r < n) ? dst_rows[r] : row0 + r;   // scattered upserts name their rows
  __nv_bfloat16* ob = out_bf16 + dst * pitch;
  float* of = out_f32 ? out_f32 + dst * pitch : nullptr;
  if (r >= n) {  // padding rows of a query batch: all zero
    for (int i = lane; i < pitch; i += 32) {
      ob[i] = __float2bfloat16_rn(0.0f);
      if (of) of[i] = 0.0f;
    }
    if (lane == 0 && norm2) norm2[dst] = 0.0;
    return;
  }
  const float* s = src + r * dim;
  double ss = 0.0;
  for (int i = lane; i < dim; i += 32) {
    float v = s[i];
    ss += (double)v * (double)v;
  }
  ss = warp_sum(ss);
  const float inv = ss > 0.0 ? (float)(1.0 / sqrt(ss)) : 0.0f;
  for (int i = lane; i < pitch; i += 32) {
    float v = i < dim ? s[i] : 0.0f;
    if (of) of[i] = v;
    ob[i] = __float2bfloat16_rn(v * inv);
  }
  if (lane == 0 && norm2) norm2[dst] = ss;
  // a NaN element makes the sum NaN, an infinite one makes it +inf (fp64 cannot overflow on fp32 squares)
  if (lane == 0 && bad_flag && !isfinite(ss)) atomicOr(bad_flag, ss != ss ? 1 : 2);
}

__device__ __forceinline__ bool scope_ok(uint32_t col, uint32_t want, int mode) {
  switch (mode) {
    case 0: return true;                      // parameter IS NULL: no constraint
    case 1: return col == 0;                  // column IS NULL
    case 2: return col != 0 && col == want;   // column = x
    default: return col == 0 || col == want;  // column IS NULL OR column = x
  }
}

// grid = (ceil(words / 8), n_filters); one thread per row, one ballot per word.
__global__ void __launch_bounds__(256) build_bitmaps_kernel(RowMetaView m, int64_t n, int words_per_group,
                                                            const DeviceFilter* __restrict__ filters,
                                                            uint32_t* __restrict__ bitmaps) {
  const DeviceFilter& f = filters[blockIdx.y];
  const int64_t row = (int64_t)blockIdx.x * 256 + threadIdx.x;
  bool ok = false;
  if (row < n) {
    const uint8_t fl = m.flags[row];
    const double n2 = m.norm2[row];  // a zero vector has no cosine: pgvector returns NaN for it
    ok = (f.workspace == kAnyWorkspaceId || m.workspace[row] == f.workspace) && n2 > 0.0 && isfinite(n2) &&
         (fl & (kRowForgotten | kRowSuperseded | kRowDeleted)) == 0 && (fl & kRowHasEmbedding) &&
         m.valid_until_us[row] > f.now_us && m.confidence[row] >= f.min_confidence &&
         scope_ok(m.user[row], f.user, f.user_mode) && scope_ok(m.agent[row], f.agent, f.agent_mode);
    if (ok && f.n_kinds > 0) {
      bool hit = false;
      const uint16_t kd = m.kind[row];
      for (int j = 0; j < f.n_kinds; j++) hit |= (kd == f.kinds[j]);
      ok = hit;
    }
    if (ok && f.n_purposes > 0) {
      bool hit = false;
      const uint16_t p = m.purpose[row];
      for (int j = 0; j < f.n_purposes; j++) hit |= (p == f.purposes[j]);
      ok = hit;
    }
  }
  const uint32_t word = __ballot_sync(0xffffffffu, ok);
  const int64_t w = row >> 5;
  if ((threadIdx.x & 31) == 0 && w < words_per_group) bitmaps[(int64_t)blockIdx.y * words_per_group + w] = word;
}

}  // namespace

int launch_ingest_rows(const float* src, int64_t n, int dim, int pitch, int64_t row0, const int64_t* dst_rows,
                       __nv_bfloat16* rows_bf16, float* rows_f32, double* norm2, int32_t* bad_flag, cudaStream_t st) {
  if (n <= 0) return 0;
  const int warps = 8;
  normalize_rows_kernel<<<(unsigned)((n + warps - 1) / warps), warps * 32, 0, st>>>(
      src, n, dim, pitch, row0, dst_rows, rows_bf16, rows_f32, norm2, n, bad_flag);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_prep_queries(const float* q, int nq, int qpad, int dim, int pitch, __nv_bfloat16* qn, double* qnorm2,
                        int32_t* bad_flag, cudaStream_t st) {
  const int warps = 8;
  static bool carveout_set = false;  // see k_post.cu: query preparation of batch i+1 runs beside scorer CTAs of batch i
  if (!carveout_set) {
    OMR_CUDA_OK(cudaFuncSetAttribute(normalize_rows_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    carveout_set = true;
  }
  normalize_rows_kernel<<<(unsigned)((qpad + warps - 1) / warps), warps * 32, 0, st>>>(
      q, nq, dim, pitch, 0, nullptr, qn, nullptr, qnorm2, qpad, bad_flag);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_build_bitmaps(RowMetaView meta, int64_t n, int words_per_group, const DeviceFilter* filters,
                         int n_filters, uint32_t* bitmaps, cudaStream_t st) {
  if (n_filters <= 0) return 0;
  dim3 grid((unsigned)((words_per_group + 7) / 8), (unsigned)n_filters);
  build_bitmaps_kernel<<<grid, 256, 0, st>>>(meta, n, words_per_group, filters, bitmaps);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
