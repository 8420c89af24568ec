// omr_exact_topk_noalloc: exact cosine top-k for ONE query without touching an allocator.
//
// Who needs this: a host process that may not grow device memory while it serves -- an
// inference server whose framework has already claimed all of HBM for its own pool, a
// request thread that must not hit cudaMalloc's device-wide synchronisation, an embedded
// caller whose memory budget was fixed at start-up.  Such a caller already owns its fp32
// vectors, its query and its result buffers on the device; the only thing the exact scan
// still needs is somewhere to put one distance per row between the scan and the selection.
// That scratch is the fixed array below, part of the library image: every GPU that loads
// the library holds it (4 MiB + 256 bytes) until the process exits.
//
// It is a variant of the engine's exact fallback (exact_scan_kernel + select_exact_kernel
// in k_select.cu): same arithmetic (fp32 inputs, fp64 products and sums, pgvector's
// 1 - clamp(dot / sqrt(na * nb))), same (distance, row) order -- rows tied at the k-th distance
// are taken smallest row number first, by an ordered scan, never by arrival order -- but on
// caller-owned rows with the row norms computed on the fly.  Nothing in the existing code paths changes.
//
// Sharing and ordering rules for users of this entry point:
//  * All calls on one GPU share the one array, and it keeps whatever the previous call left.
//    The only cell that must start at zero (the count of rankable rows) is cleared by every
//    call before its scan; distances are always fully rewritten for rows [0, n).
//  * Two calls whose GPU work overlaps (different streams, or different host threads on
//    different streams) race on the array and return garbage.  Serialise them.
//  * The first call on each device looks up the array's device address.  Make that first
//    call before starting any stream capture on that device.
#include <mutex>

#include "../../include/omnia_recall.h"
#include "common.cuh"

namespace omr {
namespace {

constexpr int64_t kMaxRows = OMR_NOALLOC_MAX_ROWS;
constexpr int kMaxK = OMR_NOALLOC_MAX_K;
constexpr size_t kHeaderBytes = 256;  // [0] unsigned long long: rows with a finite distance
constexpr size_t kScratchBytes = kHeaderBytes + (size_t)kMaxRows * sizeof(double);

// In host code this NAME is only a stand-in; the device address comes from
// cudaGetSymbolAddress below.  Device code may use the name directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One warp per row: exact distance into scratch, +inf for rows without a cosine.
__global__ void __launch_bounds__(256) noalloc_scan_kernel(const float* __restrict__ rows, int64_t n, int dim,
                                                           int64_t pitch, const float* __restrict__ q,
                                                           unsigned char* __restrict__ scratch) {
  unsigned long long* finite = reinterpret_cast<unsigned long long*>(scratch);
  double* dist = reinterpret_cast<double*>(scratch + kHeaderBytes);
  const int lane = threadIdx.x & 31;
  const int64_t row = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= n) return;
  const float* m = rows + row * pitch;
  double dot = 0.0, nm = 0.0, nq = 0.0;
  for (int i = lane; i < dim; i += 32) {
    const double a = (double)m[i], b = (double)q[i];
    dot += a * b;
    nm += a * a;
    nq += b * b;
  }
  dot = warp_sum_d(dot);
  nm = warp_sum_d(nm);
  nq = warp_sum_d(nq);
  if (lane == 0) {
    double s = dot / sqrt(nq * nm);
    s = s > 1.0 ? 1.0 : (s < -1.0 ? -1.0 : s);
    const double d = 1.0 - s;
    const bool ok = d == d;  // NaN: a zero row or a zero query never ranks
    dist[row] = ok ? d : INFINITY;
    if (ok) atomicAdd(finite, 1ull);
  }
}

// One block: the k smallest distances (their bit patterns order like unsigned integers, all
// being >= 0 or +inf) by radix select, then a bitonic sort by (distance, row).
__global__ void __launch_bounds__(1024) noalloc_select_kernel(const unsigned char* __restrict__ scratch, int64_t n,
                                                              int k, int64_t* __restrict__ out_rows,
                                                              double* __restrict__ out_dist,
                                                              int32_t* __restrict__ out_count) {
  __shared__ uint32_t hist[256];
  __shared__ unsigned long long s_prefix, s_mask;
  __shared__ uint32_t s_need, s_out;
  __shared__ double sd[kMaxK];
  __shared__ int64_t sr[kMaxK];
  const unsigned long long finite = *reinterpret_cast<const unsigned long long*>(scratch);
  const unsigned long long* d = reinterpret_cast<const unsigned long long*>(scratch + kHeaderBytes);
  const unsigned long long kInf = 0x7FF0000000000000ull;
  int P = 1;
  while (P < k) P <<= 1;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    sd[i] = INFINITY;
    sr[i] = INT64_MAX;
  }
  if (threadIdx.x == 0) {
    s_prefix = 0;
    s_mask = 0;
    s_need = (uint32_t)k;
    s_out = 0;
  }
  __syncthreads();
  unsigned long long T = kInf - 1;
  uint32_t need_eq = 0xffffffffu;
  if (finite > (unsigned long long)k) {
    for (int shift = 56; shift >= 0; shift -= 8) {
      for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
      __syncthreads();
      const unsigned long long prefix = s_prefix, mask = s_mask;
      for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
        const unsigned long long u = d[i];
        if ((u & mask) == prefix) atomicAdd(&hist[(u >> shift) & 255ull], 1u);
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        uint32_t cum = 0, need = s_need;
        int dg = 0;
        for (; dg < 255; dg++) {
          if (cum + hist[dg] >= need) break;
          cum += hist[dg];
        }
        s_need = need - cum;
        s_prefix |= (unsigned long long)dg << shift;
        s_mask |= 0xFFull << shift;
      }
      __syncthreads();
    }
    T = s_prefix;
    need_eq = s_need;
  }
  // Rows strictly nearer than T are all taken.  Of the rows AT distance T exactly `need_eq` are taken, and they
  // are the ones with the smallest row numbers: the block walks the rows in order and every tied row learns how
  // many tied rows precede it (ballot inside the warp, per-warp counts across the block), so the answer is the
  // same on every run even when exact duplicates straddle the k-th place.
  __shared__ uint32_t s_warp_ties[32];
  __shared__ uint32_t s_tie_base;
  if (threadIdx.x == 0) s_tie_base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  const bool all_rank = finite <= (unsigned long long)k;
  for (int64_t i0 = 0; i0 < n; i0 += blockDim.x) {
    const int64_t i = i0 + threadIdx.x;
    const unsigned long long u = i < n ? d[i] : kInf;
    const bool fin = u < kInf;
    const bool tie = fin && !all_rank && u == T;
    const uint32_t ballot = __ballot_sync(0xffffffffu, tie);
    if (lane == 0) s_warp_ties[wid] = __popc(ballot);
    __syncthreads();
    uint32_t before = s_tie_base + __popc(ballot & ((1u << lane) - 1u));
    for (int w = 0; w < wid; w++) before += s_warp_ties[w];
    const bool take = fin && (all_rank || u < T || (tie && before < need_eq));
    if (take) {
      const uint32_t o = atomicAdd(&s_out, 1u);
      if (o < (uint32_t)P) {
        sd[o] = __longlong_as_double((long long)u);
        sr[o] = i;
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
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      __syncthreads();
      for (int i = threadIdx.x; i < (P >> 1); i += blockDim.x) {
        const int lo = 2 * i - (i & (stride - 1)), hi = lo + stride;
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
  const int found = (int)(finite < (unsigned long long)k ? finite : (unsigned long long)k);
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    out_rows[i] = i < found ? sr[i] : -1;
    out_dist[i] = i < found ? sd[i] : INFINITY;
  }
  if (threadIdx.x == 0) *out_count = found;
}

constexpr int kMaxDevices = 64;
std::mutex g_setup_mu;
unsigned char* g_scratch_addr[kMaxDevices] = {nullptr};

// Once per device, with that device current: ask the runtime where the array lives there.
int scratch_for_current_device(unsigned char** out) {
  int dev = -1;
  OMR_CUDA_OK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail_msg(OMR_E_INVALID, "memory: gpu: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_setup_mu);
  if (!g_scratch_addr[dev]) {
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(g_noalloc_scratch)", __FILE__, __LINE__);
    if (!p) return fail_msg(OMR_E_CUDA, "memory: gpu: the static scratch array has no address on device %d", dev);
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return 0;
}

}  // namespace
}  // namespace omr

using namespace omr;

extern "C" int omr_exact_topk_noalloc(const float* d_rows, int64_t n, int32_t dim, int64_t row_pitch,
                                      const float* d_query, int32_t k, int64_t* d_out_rows, double* d_out_dist,
                                      int32_t* d_out_count, void* stream) {
  if (!d_rows || !d_query || !d_out_rows || !d_out_dist || !d_out_count)
    return fail_msg(OMR_E_INVALID, "memory: omr_exact_topk_noalloc: null argument");
  if (dim < 1 || dim > OMR_MAX_DIM)
    return fail_msg(OMR_E_DIM, "memory: embedding dimension %d outside 1..%d", dim, OMR_MAX_DIM);
  if (row_pitch < dim) return fail_msg(OMR_E_INVALID, "memory: row pitch %lld is smaller than dim %d", (long long)row_pitch, dim);
  if (n < 0) return fail_msg(OMR_E_INVALID, "memory: negative row count");
  if (n > kMaxRows)
    return fail_msg(OMR_E_CAPACITY,
                    "memory: omr_exact_topk_noalloc handles at most %lld rows per call (its scratch is a fixed "
                    "%zu-byte array), got %lld; split the scan or use omr_device_search",
                    (long long)kMaxRows, kScratchBytes, (long long)n);
  if (k < 1 || k > kMaxK)
    return fail_msg(OMR_E_CAPACITY, "memory: omr_exact_topk_noalloc: k must be in 1..%d, got %d", kMaxK, k);
  unsigned char* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  OMR_CUDA_OK(cudaMemsetAsync(scratch, 0, kHeaderBytes, st));  // the rankable-row counter starts at zero
  if (n > 0) {
    noalloc_scan_kernel<<<(unsigned)((n + 7) / 8), 256, 0, st>>>(d_rows, n, dim, row_pitch, d_query, scratch);
    OMR_CUDA_OK(cudaGetLastError());
  }
  noalloc_select_kernel<<<1, 1024, 0, st>>>(scratch, n, k, d_out_rows, d_out_dist, d_out_count);
  OMR_CUDA_OK(cudaGetLastError());
  return OMR_OK;
}
