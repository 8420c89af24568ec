// Device side of the consolidation self-join around the tensor-core kernel (k_score_umma.cu):
//   member scan     which rows take part (one row per entity of the workspace that carries a usable
//                   embedding) and their compact index list, straight from the resident metadata columns:
//                   the WHERE clause of BuildEntityDuplicateCandidatesQuery
//                   (internal/memory/consolidation/prefilter.go:118-127)
//   pair keys       exact similarity of every candidate pair (fp32 rows, fp64 sums), the `>= floor`,
//                   `e1.kind = e2.kind` and `e1.id < e2.id` rules (prefilter.go:120-128), and a 128-bit sort key
//   top-n select    `ORDER BY similarity DESC LIMIT n` (prefilter.go:129-130) by radix selection over those
//                   keys, so only n pairs ever leave the device
// Nothing here allocates; store.cu owns the buffers.
#include "kernels.cuh"

namespace omr {

namespace {

constexpr int kScanRows = 2048;  // rows per block of the member scan (8 warps x 8 words)

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ bool is_member(const RowMetaView& m, int64_t r, uint32_t ws, int require_mutable) {
  const uint8_t f = m.flags[r];
  if (m.workspace[r] != ws) return false;
  if (!(f & kRowHasEmbedding) || !(f & kRowEntityRep) || (f & kRowDeleted)) return false;
  if (require_mutable && !(f & kRowMutable)) return false;
  const double n2 = m.norm2[r];
  return n2 > 0.0 && n2 <= 1e300;  // a vector without a cosine (zero / non-finite norm) never passes `>= floor`
}

// bitmap[w] = membership of rows 32w..32w+31; block_counts[b] = members among the block's 2048 rows
__global__ void __launch_bounds__(256) member_flag_kernel(RowMetaView meta, int64_t n, uint32_t ws, int require_mutable,
                                                          uint32_t* __restrict__ bitmap, int32_t* __restrict__ block_counts) {
  __shared__ int s_count;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int mine = 0;
  for (int i = 0; i < 8; i++) {
    const int64_t word = (int64_t)blockIdx.x * (kScanRows / 32) + warp * 8 + i;
    const int64_t r = word * 32 + lane;
    const bool in = r < n && is_member(meta, r, ws, require_mutable);
    const uint32_t bits = __ballot_sync(0xffffffffu, in);
    if (lane == 0 && word * 32 < n) bitmap[word] = bits;
    mine += __popc(bits);
  }
  if (lane == 0 && mine) atomicAdd(&s_count, mine);
  __syncthreads();
  if (threadIdx.x == 0) block_counts[blockIdx.x] = s_count;
}

// exclusive scan of the block counts (one block; n_blocks is rows / 2048), total -> *total
__global__ void __launch_bounds__(1024) member_scan_kernel(const int32_t* __restrict__ block_counts, int n_blocks,
                                                           int32_t* __restrict__ block_offsets, int64_t* __restrict__ total) {
  __shared__ int s_warp[32];
  __shared__ int s_carry;
  if (threadIdx.x == 0) s_carry = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < n_blocks ? block_counts[i] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = s_carry;
    for (int w = 0; w < warp; w++) before += s_warp[w];
    if (i < n_blocks) block_offsets[i] = before + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_carry = before + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = s_carry;
}

// idx[j] = j-th member row (ascending); kind_c[j] = its kind
__global__ void __launch_bounds__(256) member_scatter_kernel(const uint32_t* __restrict__ bitmap, int64_t n,
                                                             const int32_t* __restrict__ block_offsets,
                                                             const uint16_t* __restrict__ kind, int32_t* __restrict__ idx,
                                                             uint16_t* __restrict__ kind_c) {
  __shared__ int s_warp[8];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t word0 = (int64_t)blockIdx.x * (kScanRows / 32) + warp * 8;
  const int64_t n_words = (n + 31) / 32;
  uint32_t w[8];
  int mine = 0;
  for (int i = 0; i < 8; i++) {
    w[i] = word0 + i < n_words ? bitmap[word0 + i] : 0u;
    mine += __popc(w[i]);
  }
  if (lane == 0) s_warp[warp] = mine;
  __syncthreads();
  int base = block_offsets[blockIdx.x];
  for (int v = 0; v < warp; v++) base += s_warp[v];
  for (int i = 0; i < 8; i++) {
    if ((w[i] >> lane) & 1u) {
      const int64_t r = (word0 + i) * 32 + lane;
      const int j = base + __popc(w[i] & ((1u << lane) - 1u));
      idx[j] = (int32_t)r;
      kind_c[j] = kind[r];
    }
    base += __popc(w[i]);
  }
}

// bitmap over the compact index space: rows [0, m) are members, the padding up to a row tile is not
__global__ void compact_member_bits_kernel(int64_t m, int64_t words, uint32_t* __restrict__ bits) {
  const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= words) return;
  const int64_t lo = w * 32;
  bits[w] = lo + 32 <= m ? 0xffffffffu : (lo >= m ? 0u : ((1u << (m - lo)) - 1u));
}

__device__ __forceinline__ unsigned long long ordered_double(double v) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

// One warp per candidate pair (rows of the STORE, already remapped from the compact space): exact similarity, the
// three row-pair rules, and the selection key.  A valid pair gets key = (ordered similarity, ~(canonical row,
// dupe row)), so that "largest key first" is `ORDER BY similarity DESC` with ties by (canonical, dupe) row, the
// order the tests and the host sort use; an invalid pair gets key 0.  pairs[i] is rewritten as (canonical, dupe).
__global__ void __launch_bounds__(256) pair_key_kernel(ShardView sh, PairCand* __restrict__ pairs, unsigned long long n_pairs,
                                                       const ulonglong2* __restrict__ entity_key,
                                                       const uint16_t* __restrict__ kind, int require_same_kind,
                                                       double floor_d, double* __restrict__ sim,
                                                       ulonglong2* __restrict__ keys, unsigned long long* __restrict__ n_valid) {
  const unsigned long long wid = (unsigned long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (wid >= n_pairs) return;
  const PairCand p = pairs[wid];
  const float* a = sh.rows_f32 + (int64_t)p.a * sh.pitch;
  const float* b = sh.rows_f32 + (int64_t)p.b * sh.pitch;
  double dot = 0.0;
  for (int i = lane; i < sh.dim; i += 32) dot += (double)a[i] * (double)b[i];
  dot = warp_sum_d(dot);
  if (lane != 0) return;
  double s = dot / sqrt(sh.norm2[p.a] * sh.norm2[p.b]);
  s = s > 1.0 ? 1.0 : (s < -1.0 ? -1.0 : s);
  sim[wid] = s;
  ulonglong2 key = make_ulonglong2(0ull, 0ull);
  bool ok = s >= floor_d;   // NaN fails, as in SQL
  if (ok && require_same_kind && kind[p.a] != kind[p.b]) ok = false;
  if (ok) {
    const ulonglong2 ea = entity_key[p.a], eb = entity_key[p.b];
    if (ea.x == eb.x && ea.y == eb.y) {
      ok = false;   // e1.id < e2.id can never hold for one entity
    } else {
      const bool a_first = ea.x < eb.x || (ea.x == eb.x && ea.y < eb.y);
      const int32_t c = a_first ? p.a : p.b, d = a_first ? p.b : p.a;
      pairs[wid] = PairCand{c, d};
      key.x = ordered_double(s);
      key.y = ~(((unsigned long long)(uint32_t)c << 32) | (unsigned long long)(uint32_t)d);
      atomicAdd(n_valid, 1ull);
    }
  }
  keys[wid] = key;
}

// ---- radix selection of the n largest 128-bit keys ------------------------------------------
struct RadixSel {
  unsigned long long prefix_hi, prefix_lo, mask_hi, mask_lo;
  unsigned long long need;
  unsigned int hist[256];
  unsigned int blocks_done;
};

// pass p = 0..15 looks at byte (15 - p) of the key, counting from the least significant byte of `lo`
__global__ void __launch_bounds__(256) radix_pass_kernel(const ulonglong2* __restrict__ keys, unsigned long long n,
                                                         RadixSel* __restrict__ st, int pass) {
  __shared__ unsigned int sh[256];
  __shared__ bool s_last;
  sh[threadIdx.x] = 0u;
  __syncthreads();
  const unsigned long long phi = st->prefix_hi, plo = st->prefix_lo, mhi = st->mask_hi, mlo = st->mask_lo;
  const bool in_hi = pass < 8;
  const int shift = in_hi ? 56 - 8 * pass : 56 - 8 * (pass - 8);
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const ulonglong2 k = keys[i];
    if (k.x != 0ull && (k.x & mhi) == phi && (k.y & mlo) == plo)
      atomicAdd(&sh[(unsigned)(((in_hi ? k.x : k.y) >> shift) & 255ull)], 1u);
  }
  __syncthreads();
  if (sh[threadIdx.x]) atomicAdd(&st->hist[threadIdx.x], sh[threadIdx.x]);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(&st->blocks_done, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x == 0) {
    // largest first: the digit at which the running count reaches `need`
    unsigned long long cum = 0, need = st->need;
    int d = 255;
    for (; d > 0; d--) {
      const unsigned int h = *reinterpret_cast<volatile unsigned int*>(&st->hist[d]);
      if (cum + h >= need) break;
      cum += h;
    }
    st->need = need > cum ? need - cum : 1ull;
    if (in_hi) {
      st->prefix_hi = phi | ((unsigned long long)d << shift);
      st->mask_hi = mhi | (0xFFull << shift);
    } else {
      st->prefix_lo = plo | ((unsigned long long)d << shift);
      st->mask_lo = mlo | (0xFFull << shift);
    }
    st->blocks_done = 0u;
  }
  __syncthreads();
  st->hist[threadIdx.x] = 0u;
}

// every valid pair whose key is at least the selected one (all valid pairs when limit == 0)
__global__ void __launch_bounds__(256) pair_compact_kernel(const PairCand* __restrict__ pairs, const double* __restrict__ sim,
                                                           const ulonglong2* __restrict__ keys, unsigned long long n,
                                                           const RadixSel* __restrict__ st, int use_threshold,
                                                           unsigned long long out_cap, PairCand* __restrict__ out_pairs,
                                                           double* __restrict__ out_sim, unsigned long long* __restrict__ out_count) {
  const unsigned long long thi = use_threshold ? st->prefix_hi : 0ull, tlo = use_threshold ? st->prefix_lo : 0ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x) {
    const ulonglong2 k = keys[i];
    if (k.x != 0ull && (k.x > thi || (k.x == thi && k.y >= tlo))) {
      const unsigned long long pos = atomicAdd(out_count, 1ull);
      if (pos < out_cap) {
        out_pairs[pos] = pairs[i];
        out_sim[pos] = sim[i];
      }
    }
  }
}

}  // namespace

int join_scan_blocks(int64_t n) { return (int)((n + kScanRows - 1) / kScanRows); }

int launch_join_member_scan(const RowMetaView& meta, int64_t n, uint32_t workspace, int require_mutable, uint32_t* bitmap,
                            int32_t* block_counts, int32_t* block_offsets, int64_t* d_total, cudaStream_t st) {
  if (n <= 0) return 0;
  const int blocks = join_scan_blocks(n);
  member_flag_kernel<<<blocks, 256, 0, st>>>(meta, n, workspace, require_mutable, bitmap, block_counts);
  OMR_CUDA_OK(cudaGetLastError());
  member_scan_kernel<<<1, 1024, 0, st>>>(block_counts, blocks, block_offsets, d_total);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_join_member_scatter(const uint32_t* bitmap, int64_t n, const int32_t* block_offsets, const uint16_t* kind,
                               int32_t* idx, uint16_t* kind_c, cudaStream_t st) {
  if (n <= 0) return 0;
  member_scatter_kernel<<<join_scan_blocks(n), 256, 0, st>>>(bitmap, n, block_offsets, kind, idx, kind_c);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

int launch_compact_member_bits(int64_t m, int64_t words, uint32_t* bits, cudaStream_t st) {
  if (words <= 0) return 0;
  compact_member_bits_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(m, words, bits);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

size_t join_select_state_bytes() { return sizeof(RadixSel); }

int launch_pair_keys(const ShardView& sh, PairCand* pairs, unsigned long long n_pairs, const void* entity_key,
                     const uint16_t* kind, int require_same_kind, double floor_d, double* sim, void* keys,
                     unsigned long long* n_valid, cudaStream_t st) {
  if (n_pairs == 0) return 0;
  pair_key_kernel<<<(unsigned)((n_pairs + 7) / 8), 256, 0, st>>>(sh, pairs, n_pairs,
                                                               reinterpret_cast<const ulonglong2*>(entity_key), kind,
                                                               require_same_kind, floor_d, sim,
                                                               reinterpret_cast<ulonglong2*>(keys), n_valid);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

// Selects the `limit` largest keys (limit <= 0: every valid pair) and compacts the matching pairs into
// out_pairs / out_sim (at most out_cap are written; *out_count receives how many qualified).  `state` is
// join_select_state_bytes() of device scratch; out_count must be zero on entry.
int launch_pair_select(const PairCand* pairs, const double* sim, const void* keys, unsigned long long n_pairs, int limit,
                       void* state, unsigned long long out_cap, PairCand* out_pairs, double* out_sim,
                       unsigned long long* out_count, int num_sms, cudaStream_t st) {
  if (n_pairs == 0) return 0;
  const ulonglong2* k = reinterpret_cast<const ulonglong2*>(keys);
  RadixSel* rs = reinterpret_cast<RadixSel*>(state);
  const unsigned grid = (unsigned)std::min<unsigned long long>((n_pairs + 255) / 256, (unsigned long long)num_sms * 8);
  if (limit > 0) {
    RadixSel init{};
    init.need = (unsigned long long)limit;
    OMR_CUDA_OK(cudaMemcpyAsync(rs, &init, sizeof init, cudaMemcpyHostToDevice, st));
    OMR_CUDA_OK(cudaStreamSynchronize(st));   // `init` is a stack object
    for (int pass = 0; pass < 16; pass++) {
      radix_pass_kernel<<<grid, 256, 0, st>>>(k, n_pairs, rs, pass);
      OMR_CUDA_OK(cudaGetLastError());
    }
  }
  pair_compact_kernel<<<grid, 256, 0, st>>>(pairs, sim, k, n_pairs, rs, limit > 0 ? 1 : 0, out_cap, out_pairs, out_sim,
                                            out_count);
  OMR_CUDA_OK(cudaGetLastError());
  return 0;
}

}  // namespace omr
