// Candidate exchange between the row shards of one box, fused over NVLink peer memory.
//
// every GPU holds a row shard, answers a query batch with its exact local
// top-k, and the per-shard answers are merged into the global top-k.  The baseline exchange is an
// NCCL allgather issued by the host (omnia_b200/dist.py).  Here the exchange is two small kernels
// with no library and no host in between:
//
//   exchange_scatter_kernel  (rank r, right behind its post_kernel on the same stream) stores r's
//                            [2][nq][k] answer into slot region r of EVERY rank's exchange buffer
//                            -- plain st.global on peer pointers mapped through CUDA IPC, i.e.
//                            NVLink / NVSwitch writes -- then publishes a per-(slot, rank) epoch flag
//                            with system-scope release semantics.
//   exchange_wait_kernel     (every rank, one warp) waits with system-scope acquire loads until all
//                            `world` flags of the slot carry the batch's epoch;
//   exchange_merge_kernel    then merges G*k -> k per query, reading the slot in place.
//
// Ranks run on DIFFERENT GPUs, so a merge kernel spinning on a peer's flag cannot starve that peer.
// The spin is bounded (about four seconds of clock64): on expiry the kernel raises an error word the
// host checks, instead of hanging the GPU.
#include "kernels.cuh"

namespace omr {

namespace {

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// grid = (blocks_per_peer, world).  Block (b, p) copies a slice of the local answer into peer p.
__global__ void __launch_bounds__(256) exchange_scatter_kernel(ExchangeView x, const int64_t* __restrict__ local,
                                                               int64_t n_elems, int slot, unsigned long long epoch,
                                                               const int32_t* __restrict__ fail_count,
                                                               unsigned int* __restrict__ done_blocks) {
  const int p = blockIdx.y;
  uint8_t* peer = reinterpret_cast<uint8_t*>(x.peers[p]);
  int64_t* dst = reinterpret_cast<int64_t*>(peer + x.data_offset(slot, x.rank));
  const int64_t n2 = n_elems >> 1;  // 16-byte vectors
  const longlong2* s2 = reinterpret_cast<const longlong2*>(local);
  longlong2* d2 = reinterpret_cast<longlong2*>(dst);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (int64_t)gridDim.x * blockDim.x)
    d2[i] = s2[i];
  if ((n_elems & 1) && blockIdx.x == 0 && threadIdx.x == 0) dst[n_elems - 1] = local[n_elems - 1];
  // the last block to finish for peer p publishes the flag (after every block's stores are visible)
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(&done_blocks[p], 1u);
    if (prev == gridDim.x - 1) {
      done_blocks[p] = 0;  // ready for the next launch on this stream
      // a batch whose certificate failed somewhere is published by the host after its exact fallback
      if (fail_count == nullptr || *fail_count == 0) {
        __threadfence_system();
        st_release_sys(reinterpret_cast<unsigned long long*>(peer + x.flag_offset(slot, x.rank)), epoch);
      }
    }
  }
}

// One warp: wait until every rank's flag of the slot carries this batch's epoch.  Kept apart from the
// merge so that only `world` threads poll instead of `world` threads in each of nq blocks.
__global__ void __launch_bounds__(32) exchange_wait_kernel(ExchangeView x, int slot, unsigned long long epoch,
                                                           unsigned int* __restrict__ error_word) {
  const uint8_t* own = reinterpret_cast<const uint8_t*>(x.peers[x.rank]);
  if ((int)threadIdx.x < x.world) {
    const unsigned long long* f = reinterpret_cast<const unsigned long long*>(own + x.flag_offset(slot, threadIdx.x));
    const long long t0 = clock64();
    while (ld_acquire_sys(f) != epoch) {
      if (clock64() - t0 > 8000000000ll) {  // ~4 s: a peer died or the callers disagree about the batch order
        atomicExch(error_word, 1u);
        break;
      }
      __nanosleep(100);
    }
  }
}

// One block per query: the same bitonic merge as merge_topk_kernel, reading the slot in place.
__global__ void __launch_bounds__(256) exchange_merge_kernel(ExchangeView x, int slot, unsigned long long epoch,
                                                             int nq, int k, int P, int64_t* __restrict__ out_rows,
                                                             double* __restrict__ out_dist,
                                                             int32_t* __restrict__ out_count,
                                                             unsigned int* __restrict__ error_word) {
  extern __shared__ __align__(16) uint8_t esm[];
  double* sd = reinterpret_cast<double*>(esm);
  int64_t* sr = reinterpret_cast<int64_t*>(sd + P);
  const uint8_t* own = reinterpret_cast<const uint8_t*>(x.peers[x.rank]);
  (void)epoch;
  const int q = blockIdx.x;
  const int G = x.world;
  // A peer never published this batch (the wait kernel gave up): the slot holds stale or partial data.  The answer
  // is marked unusable (count -1, empty rows) instead of merged; the host raises the error with the next call.
  if (*reinterpret_cast<volatile unsigned int*>(error_word) != 0u) {
    for (int i = threadIdx.x; i < k; i += blockDim.x) {
      out_rows[(int64_t)q * k + i] = -1;
      out_dist[(int64_t)q * k + i] = INFINITY;
    }
    if (threadIdx.x == 0) out_count[q] = -1;
    return;
  }
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    int64_t r = -1;
    double d = INFINITY;
    if (i < G * k) {
      const int g = i / k, j = i % k;
      const int64_t* base = reinterpret_cast<const int64_t*>(own + x.data_offset(slot, g));
      r = base[(int64_t)q * k + j];
      d = __longlong_as_double(base[(int64_t)nq * k + (int64_t)q * k + j]);
    }
    sd[i] = r >= 0 ? d : INFINITY;
    sr[i] = r >= 0 ? r : INT64_MAX;
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
  for (int i = threadIdx.x; i < k; i += blockDim.x) {
    const bool ok = i < P && sr[i] != INT64_MAX;
    out_rows[(int64_t)q * k + i] = ok ? sr[i] : -1;
    out_dist[(int64_t)q * k + i] = ok ? sd[i] : INFINITY;
  }
  if (threadIdx.x == 0) {
    int cnt = 0;
    const int lim = P < k ? P : k;
    while (cnt < lim && sr[cnt] != INT64_MAX) cnt++;
    out_count[q] = cnt;
  }
}

}  // namespace

int launch_exchange_scatter(const ExchangeView& x, const int64_t* local, int nq, int k, int slot,
                            unsigned long long epoch, const int32_t* fail_count, unsigned int* done_blocks,
                            cudaStream_t st) {
  static bool carveout_set = false;  // co-resident with scorer CTAs of the next batch (see k_post.cu)
  if (!carveout_set) {
    OMR_CUDA_OK(cudaFuncSetAttribute(exchange_scatter_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    carveout_set = true;
  }
  dim3 grid(4, (unsigned)x.world);
  exchange_scatter_kernel<<<grid, 256, 0, st>>>(x, local, (int64_t)2 * nq * k, slot, epoch, fail_count, done_blocks);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_exchange_merge(const ExchangeView& x, int slot, unsigned long long epoch, int nq, int k,
                          int64_t* out_rows, double* out_dist, int32_t* out_count, unsigned int* error_word,
                          cudaStream_t st) {
  int P = 1;
  while (P < x.world * k) P <<= 1;
  const size_t smem = (size_t)P * 16;
  if (smem > 200 * 1024) return fail_msg(-2, "merge of %d shards x k=%d exceeds the 12800-entry limit", x.world, k);
  static bool carveout_set = false;
  if (!carveout_set) {
    OMR_CUDA_OK(cudaFuncSetAttribute(exchange_merge_kernel, cudaFuncAttributePreferredSharedMemoryCarveout,
                                     cudaSharedmemCarveoutMaxShared));
    carveout_set = true;
  }
  if (smem > 32 * 1024)  // static shared memory counts against the 48 KB default too
    OMR_CUDA_OK(cudaFuncSetAttribute(exchange_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  exchange_wait_kernel<<<1, 32, 0, st>>>(x, slot, epoch, error_word);
  OMR_CUDA_OK(cudaGetLastError());
  exchange_merge_kernel<<<nq, 256, smem, st>>>(x, slot, epoch, nq, k, P, out_rows, out_dist, out_count, error_word);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
