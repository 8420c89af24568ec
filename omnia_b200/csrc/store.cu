// Host side of the recall engine: the resident store and the C ABI above it.
//
// The reference keeps embeddings in Postgres (`memory_observations.embedding`,
// internal/memory/postgres/embedding_schema.go:103-111) and answers recall with SQL
// (internal/memory/store_read.go:243-306, internal/memory/store.go:240-272).  Here the
// same rows live in HBM: an fp32 master matrix (exact rescoring), a unit-normalised bf16
// copy (what the scorers stream), fp64 squared norms and the metadata columns the WHERE
// clause reads, all struct-of-arrays.  The host keeps uuid <-> row maps and mirrors of
// the metadata so that row-state edits are one small cudaMemcpy.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/omnia_recall.h"
#include "kernels.cuh"

namespace omr {

static thread_local std::string g_err;

int fail_msg(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

int fail_cuda(cudaError_t e, const char* what, const char* file, int line) {
  return fail_msg(OMR_E_CUDA, "memory: gpu: %s failed: %s (%s:%d)", what, cudaGetErrorString(e), file, line);
}

extern bool g_force_ldg_stream;
int launch_iota(int32_t* p, int n, cudaStream_t st);
int launch_fill_empty(int64_t* rows, double* dist, int32_t* count, int nq, int k, cudaStream_t st);
int launch_pair_remap(PairCand* pairs, unsigned long long n, const int32_t* remap, cudaStream_t st);
// k_join.cu
int join_scan_blocks(int64_t n);
int launch_join_member_scan(const RowMetaView& meta, int64_t n, uint32_t workspace, int require_mutable, uint32_t* bitmap,
                            int32_t* block_counts, int32_t* block_offsets, int64_t* d_total, cudaStream_t st);
int launch_join_member_scatter(const uint32_t* bitmap, int64_t n, const int32_t* block_offsets, const uint16_t* kind,
                               int32_t* idx, uint16_t* kind_c, cudaStream_t st);
int launch_compact_member_bits(int64_t m, int64_t words, uint32_t* bits, cudaStream_t st);
size_t join_select_state_bytes();
int launch_pair_keys(const ShardView& sh, PairCand* pairs, unsigned long long n_pairs, const void* entity_key,
                     const uint16_t* kind, int require_same_kind, double floor_d, double* sim, void* keys,
                     unsigned long long* n_valid, cudaStream_t st);
int launch_pair_select(const PairCand* pairs, const double* sim, const void* keys, unsigned long long n_pairs, int limit,
                       void* state, unsigned long long out_cap, PairCand* out_pairs, double* out_sim,
                       unsigned long long* out_count, int num_sms, cudaStream_t st);

namespace {

struct Key128 {
  uint64_t a, b;
  bool operator==(const Key128& o) const { return a == o.a && b == o.b; }
};
struct Key128Hash {
  size_t operator()(const Key128& k) const { return (size_t)(k.a * 0x9E3779B97F4A7C15ull ^ (k.b + 0x7F4A7C15ull)); }
};
Key128 key_of(const uint8_t* p) {
  Key128 k;
  memcpy(&k.a, p, 8);
  memcpy(&k.b, p + 8, 8);
  return k;
}
void key_to(const Key128& k, uint8_t* p) {
  memcpy(p, &k.a, 8);
  memcpy(p + 8, &k.b, 8);
}
Key128 key_from_index(int64_t i) {  // default ids: big-endian row number in the low bytes
  uint8_t b[16] = {0};
  for (int j = 0; j < 8; j++) b[15 - j] = (uint8_t)((uint64_t)i >> (8 * j));
  return key_of(b);
}

struct Interner {
  std::unordered_map<std::string, uint32_t> ids;
  std::vector<std::string> names;
  uint32_t base;
  explicit Interner(uint32_t b) : base(b) {}
  uint32_t intern(const std::string& s) {
    auto it = ids.find(s);
    if (it != ids.end()) return it->second;
    uint32_t id = base + (uint32_t)names.size();
    names.push_back(s);
    ids.emplace(s, id);
    return id;
  }
  bool lookup(const char* s, uint32_t* out) const {
    auto it = ids.find(s);
    if (it == ids.end()) return false;
    *out = it->second;
    return true;
  }
};

constexpr uint32_t kNoSuchId = 0xFFFFFFFEu;  // matches no row
constexpr uint32_t kAnyWorkspace = kAnyWorkspaceId;
constexpr int kExtraSlots = 1024;

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  int reserve(size_t want) {
    if (want <= n) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
    OMR_CUDA_OK(cudaMalloc(&p, want * sizeof(T)));
    n = want;
    return 0;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
};

template <typename T>
struct PinBuf {
  T* p = nullptr;
  size_t n = 0;
  int reserve(size_t want) {
    if (want <= n) return 0;
    if (p) cudaFreeHost(p);
    p = nullptr;
    n = 0;
    OMR_CUDA_OK(cudaHostAlloc(&p, want * sizeof(T), cudaHostAllocDefault));
    n = want;
    return 0;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    n = 0;
  }
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int pow2ceil_i(int v) {
  int p = 1;
  while (p < v) p <<= 1;
  return p;
}

}  // namespace
}  // namespace omr

using namespace omr;

struct omr_store {
  omr_config cfg{};
  int dim = 0, pitch = 0, num_sms = 148;
  int64_t cap = 0, n = 0;
  float margin = 4.5e-3f;
  std::mutex mu;
  cudaStream_t stream = nullptr;
  EncodeTiledFn encode = nullptr;

  // resident matrices and metadata (device)
  __nv_bfloat16* d_bf16 = nullptr;
  float* d_f32 = nullptr;
  double* d_norm2 = nullptr;
  uint32_t *d_ws = nullptr, *d_user = nullptr, *d_agent = nullptr;
  uint16_t *d_kind = nullptr, *d_purpose = nullptr;
  float* d_conf = nullptr;
  uint8_t* d_flags = nullptr;
  int64_t* d_valid = nullptr;
  unsigned long long* d_entity_key = nullptr;   // [cap][2]: the row's entity uuid as two big-endian words (memcmp order)
  CUtensorMap tmap_rows_b{}, tmap_rows_a{};

  // host mirrors and identity maps
  std::vector<uint32_t> h_ws, h_user, h_agent, h_entity;
  std::vector<uint16_t> h_kind, h_purpose;
  std::vector<float> h_conf;
  std::vector<uint8_t> h_flags;
  std::vector<int64_t> h_valid;
  std::vector<Key128> obs_id;
  std::vector<Key128> entity_id;                 // by entity index
  std::vector<int64_t> entity_latest;            // entity index -> representative row
  std::vector<std::vector<int64_t>> entity_rows; // entity index -> rows
  std::unordered_map<Key128, int64_t, Key128Hash> obs_row;
  std::unordered_map<Key128, uint32_t, Key128Hash> entity_idx;
  Interner ws_names{1}, user_names{1}, agent_names{1}, kind_names{0}, purpose_names{0};

  // Per-search scratch (grown on demand, reused).  A Lane is everything one batch in flight needs to
  // itself: lane 0 serves the synchronous entry points, lanes 1 and 2 the begin/finish pipeline.
  struct Lane {
    DevBuf<__nv_bfloat16> qn;
    DevBuf<double> qnorm2;
    DevBuf<int32_t> cand, fail_list, counters, extra_cnt;
    // emission mode (tensor-core scorers, k <= 256): per-query thresholds, counts and candidate lists
    DevBuf<int32_t> ethr, seed_keys;
    DevBuf<unsigned int> ecount;
    DevBuf<unsigned long long> elist;
    bool elist_dirty = true;   // a batch was enqueued past the scorer without its post kernel: lists may hold entries
    DevBuf<long long> timing;
    PinBuf<int32_t> pin;
    CUtensorMap tmap_q{};
    size_t tmap_q_rows = 0;
    const void* tmap_q_ptr = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_in = nullptr, ev_done = nullptr, ev0 = nullptr, ev1 = nullptr;
    // The finishing kernel of a batch runs on a LOW-priority stream of its own (ordered by events): its thousand
    // small blocks must not take the SMs ahead of the next batch's scorer CTAs, which need a whole SM's shared
    // memory each and fit beside at most one of them.
    cudaStream_t pstream = nullptr;
    cudaEvent_t ev_scored = nullptr, ev_post = nullptr;
    bool busy = false, immediate = false, timed = false, phased = false;
    // the batch in flight
    const float* d_q = nullptr;
    const uint32_t* d_bitmaps = nullptr;
    const int32_t* d_qgroup = nullptr;
    int nq = 0, k = 0, K2 = 0, K2x = 0, qpad = 0, nfail_known = -1;
    cudaStream_t caller = nullptr;
    FinalizeArgs fa;
    // fused exchange: this batch's local answer [2][nq][k] + counts, its slot and epoch
    DevBuf<int64_t> xlocal;
    DevBuf<int32_t> xcount;
    int xslot = -1, xnq = 0, xk = 0, nfail_seen = 0;
    unsigned long long xepoch = 0;
    // host-buffer batches in flight (omr_search_rows_begin/_finish)
    DevBuf<float> hq;
    DevBuf<int32_t> hgroup, hcount;
    DevBuf<double> hmaxd, hdist;
    DevBuf<int64_t> hrows;
    DevBuf<uint32_t> hbitmaps;
    DevBuf<DeviceFilter> hfilters;
    PinBuf<uint8_t> hin, hout;
    int hnq = 0, hk = 0;
    bool host_batch = false;
  };
  static constexpr int kLanes = 5;   // lane 0: synchronous calls; lanes 1..4: batches in flight
  Lane lanes[kLanes];
  DevBuf<double> w_sel_dist, w_exact, w_maxdist, w_out_dist, w_pair_sim;
  DevBuf<int32_t> w_sel_rows, w_extra_items, w_qgroup, w_out_count, w_idx;
  DevBuf<float> w_sel_bound, w_queries, w_stage;
  DevBuf<uint32_t> w_bitmaps, w_live, w_member;
  DevBuf<int64_t> w_out_rows;
  DevBuf<DeviceFilter> w_filters;
  // ---- self-join (its own stream and scratch: recall keeps running while a join is in flight) ----
  DevBuf<PairCand> w_pairs, w_out_pairs;
  DevBuf<unsigned long long> w_pair_count;   // [0] candidates, [1] valid pairs, [2] selected pairs
  DevBuf<__nv_bfloat16> w_gather;
  DevBuf<int32_t> w_jcounts, w_joffsets, w_run_cut;
  DevBuf<int64_t> w_jtotal;
  DevBuf<uint16_t> w_kind_c;
  DevBuf<uint32_t> w_jbits, w_ghist;
  DevBuf<unsigned long long> w_keys;
  DevBuf<uint8_t> w_jstate;
  DevBuf<double> w_out_sim;
  PinBuf<unsigned long long> jpin;
  cudaStream_t jstream = nullptr;
  cudaEvent_t jev0 = nullptr, jev1 = nullptr, jchunk[2] = {nullptr, nullptr};
  // Lock order: join_gate, then mu.  A join holds join_gate from start to end and mu only while it reads or
  // writes host state; every call that changes rows takes join_gate first, so it waits for a running join, while
  // recall (mu only) proceeds beside it.
  std::mutex join_gate;
  int selfjoin(const omr_selfjoin_opts* o, std::vector<PairCand>* pairs, std::vector<double>* sims);
  PinBuf<uint8_t> pin_in, pin_out;
  bool live_dirty = true;
  bool post_timing = false;  // OMR_POST_TIMING=1: per-phase clock stamps of the fused post kernel
  bool legacy_post = false;  // OMR_LEGACY_POST=1: use the unfused select/rescore/finalize chain
  bool use_pdl = true;       // OMR_PDL=0: plain stream order between seed scorer, seed thresholds and scorer

  std::atomic<int64_t> st_launches{0}, st_searches{0}, st_queries{0}, st_rescued{0}, st_fallback{0}, st_listed{0}, st_join_candidates{0};
  // Coalescing of concurrent host calls (dedup-on-write micro-batching): while one
  // search runs, the calls that arrive queue up and the next leader takes them all as ONE batch.
  struct HostReq {
    const float* q; int nq; const omr_filter* filters; int n_filters; const int32_t* qf; int k;
    omr_hit* hits; int32_t* counts; int64_t* rows; double* dist;
    int rc = 0; bool done = false; std::string err;
  };
  std::mutex bq_mu;
  std::condition_variable bq_cv;
  std::deque<HostReq*> bq;
  bool bq_leader = false;
  std::atomic<int64_t> st_host_calls{0}, st_host_batches{0};
  // fused exchange state (omr_exchange_*)
  ExchangeView xv;
  void* xbuf = nullptr;               // this rank's exchange buffer (cudaMalloc, exported through CUDA IPC)
  std::vector<void*> xopened;         // peer mappings to close
  std::vector<unsigned long long> xslot_epoch;
  int xnq_max = 0, xk_max = 0;
  DevBuf<unsigned int> xwords;        // [0] error word, then 16 scatter block counters per lane
  DevBuf<int32_t> counters_init;      // what a lane's counters start a batch with: zeros, and ~0 where a time stamp takes a minimum
  PinBuf<unsigned int> xerr_pin;
  int submit(HostReq& r);
  void run_batch(std::vector<HostReq*>& batch);
  // optional device timing of the dominant kernel (the approximate scorer), for bench.py's roofline
  bool profiling = false;
  double scorer_ms_sum = 0.0, phase_ms_sum[3] = {0.0, 0.0, 0.0}, kernel_ns_sum = 0.0, kernel_start_skew_ns = 0.0,
         kernel_end_skew_ns = 0.0;
  int64_t scorer_timed = 0, kernel_ns_n = 0;
  unsigned long long last_main_end_ns = 0;
  double gap_ns_sum = 0.0, post_ns_sum = 0.0, post_lag_ns_sum = 0.0;
  int last_path = 0, last_S = 0, last_K2 = 0;

  int64_t words() const { return (cap / kTileRows) * 8; }
  ShardView view() const { return ShardView{d_bf16, d_f32, d_norm2, n, dim, pitch}; }
  RowMetaView meta() const { return RowMetaView{d_ws, d_user, d_agent, d_kind, d_purpose, d_conf, d_flags, d_valid, d_norm2}; }

  int make_tmap(CUtensorMap* out, const void* ptr, uint64_t rows, uint32_t box_rows) {
    cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
    cuuint32_t box[2] = {(cuuint32_t)kBlockK, box_rows};
    cuuint32_t es[2] = {1, 1};
    CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, es,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail_msg(OMR_E_CUDA, "memory: gpu: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
  }

  int open();
  void destroy();
  int append(int64_t count, const float* vecs, int64_t row0, const int64_t* dst = nullptr);
  void release_row(int64_t r);
  int64_t recycle_row(std::vector<int64_t>* touched);
  int compact(int64_t* reclaimed);
  std::vector<int64_t> free_rows;   // hard-deleted slots waiting for a new tenant (entries may be stale: see recycle_row)
  DevBuf<int32_t> w_bad, w_cidx;
  DevBuf<int64_t> w_dst;
  int push_meta(int64_t row0, int64_t count);
  int push_flag(int64_t row);
  int to_device_filter(const omr_filter& f, DeviceFilter* out, double* maxd);
  int ensure_live(cudaStream_t st);
  int device_search(const float* d_q, int nq, const uint32_t* d_bitmaps, const int32_t* d_qgroup,
                    const double* d_maxdist, int k, int64_t* d_rows, double* d_dist, int32_t* d_count, cudaStream_t st);
  int enqueue(Lane& L, const float* d_q, int nq, const uint32_t* d_bitmaps, const int32_t* d_qgroup,
              const double* d_maxdist, int k, int64_t* d_rows, double* d_dist, int32_t* d_count, cudaStream_t st,
              bool* fused_out);
  int complete(Lane& L, cudaStream_t st);
  int prepare_emit(Lane& L, int qpad, int k, int n_tiles, cudaStream_t st, EmitView* ev, int* seed_n, int64_t* seed_kpq);
  int drain();
  int host_search(const float* queries, int nq, const omr_filter* filters, int n_filters, const int32_t* query_filter,
                  int k, omr_hit* hits, int32_t* counts, int64_t* soa_rows = nullptr, double* soa_dist = nullptr);
  void fill_hit(omr_hit* h, int64_t grow, double dist) const;
};

int omr_store::open() {
  OMR_CUDA_OK(cudaSetDevice(cfg.device));
  cudaDeviceProp prop;
  OMR_CUDA_OK(cudaGetDeviceProperties(&prop, cfg.device));
  if (prop.major != 10)
    return fail_msg(OMR_E_CUDA, "memory: gpu: device %d is sm_%d%d; this library is built for sm_100a only", cfg.device,
                    prop.major, prop.minor);
  num_sms = prop.multiProcessorCount;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  OMR_CUDA_OK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess)
    return fail_msg(OMR_E_CUDA, "memory: gpu: cuTensorMapEncodeTiled not available from the driver");
  encode = (EncodeTiledFn)fn;
  OMR_CUDA_OK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
  const size_t c = (size_t)cap;
  OMR_CUDA_OK(cudaMalloc(&d_bf16, c * pitch * sizeof(__nv_bfloat16)));
  OMR_CUDA_OK(cudaMalloc(&d_f32, c * pitch * sizeof(float)));
  OMR_CUDA_OK(cudaMalloc(&d_norm2, c * sizeof(double)));
  OMR_CUDA_OK(cudaMalloc(&d_ws, c * 4));
  OMR_CUDA_OK(cudaMalloc(&d_user, c * 4));
  OMR_CUDA_OK(cudaMalloc(&d_agent, c * 4));
  OMR_CUDA_OK(cudaMalloc(&d_kind, c * 2));
  OMR_CUDA_OK(cudaMalloc(&d_purpose, c * 2));
  OMR_CUDA_OK(cudaMalloc(&d_conf, c * 4));
  OMR_CUDA_OK(cudaMalloc(&d_flags, c));
  OMR_CUDA_OK(cudaMalloc(&d_valid, c * 8));
  OMR_CUDA_OK(cudaMalloc(&d_entity_key, c * 16));
  OMR_CUDA_OK(cudaStreamCreateWithFlags(&jstream, cudaStreamNonBlocking));
  OMR_CUDA_OK(cudaEventCreate(&jev0));
  OMR_CUDA_OK(cudaEventCreate(&jev1));
  for (cudaEvent_t& e : jchunk) OMR_CUDA_OK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  OMR_CUDA_OK(cudaMemset(d_bf16, 0, c * pitch * sizeof(__nv_bfloat16)));
  OMR_CUDA_OK(cudaMemset(d_f32, 0, c * pitch * sizeof(float)));
  OMR_CUDA_OK(cudaMemset(d_norm2, 0, c * sizeof(double)));
  OMR_CUDA_OK(cudaMemset(d_flags, 0, c));
  OMR_CUDA_OK(cudaMemset(d_ws, 0, c * 4));
  int rc = make_tmap(&tmap_rows_b, d_bf16, (uint64_t)cap, kTileRows);
  if (rc) return rc;
  rc = make_tmap(&tmap_rows_a, d_bf16, (uint64_t)cap, kQueryTile);
  if (rc) return rc;
  {
    // [0..3] counters = 0; then six 64-bit stamps: scorer earliest start (min), latest end, latest start, earliest end
    // (min), post earliest start (min), post latest end
    if ((rc = counters_init.reserve(16))) return rc;
    unsigned long long t[8] = {0ull, 0ull, ~0ull, 0ull, 0ull, ~0ull, ~0ull, 0ull};
    OMR_CUDA_OK(cudaMemcpy(counters_init.p, t, sizeof(t), cudaMemcpyHostToDevice));
  }
  for (Lane& L : lanes) {
    int prio_lo = 0, prio_hi = 0;
    OMR_CUDA_OK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));   // numerically lower = higher priority
    OMR_CUDA_OK(cudaStreamCreateWithPriority(&L.stream, cudaStreamNonBlocking, prio_hi));
    OMR_CUDA_OK(cudaStreamCreateWithPriority(&L.pstream, cudaStreamNonBlocking, prio_lo));
    OMR_CUDA_OK(cudaEventCreateWithFlags(&L.ev_scored, cudaEventDisableTiming));
    OMR_CUDA_OK(cudaEventCreateWithFlags(&L.ev_post, cudaEventDisableTiming));
    OMR_CUDA_OK(cudaEventCreateWithFlags(&L.ev_in, cudaEventDisableTiming));
    OMR_CUDA_OK(cudaEventCreateWithFlags(&L.ev_done, cudaEventDisableTiming));
    OMR_CUDA_OK(cudaEventCreate(&L.ev0));
    OMR_CUDA_OK(cudaEventCreate(&L.ev1));
    if ((rc = L.pin.reserve(16))) return rc;
    if ((rc = L.counters.reserve(16))) return rc;   // 4 counters, then (profiling) 6 x 64-bit kernel time stamps
  }
  return 0;
}

void omr_store::destroy() {
  cudaSetDevice(cfg.device);
  if (stream) cudaStreamSynchronize(stream);
  if (jstream) cudaStreamSynchronize(jstream);
  void* ptrs[] = {d_bf16, d_f32, d_norm2, d_ws, d_user, d_agent, d_kind, d_purpose, d_conf, d_flags, d_valid, d_entity_key};
  for (void* p : ptrs)
    if (p) cudaFree(p);
  for (Lane& L : lanes) {
    if (L.stream) cudaStreamSynchronize(L.stream);
    L.xlocal.release(); L.xcount.release();
    L.hq.release(); L.hgroup.release(); L.hcount.release(); L.hmaxd.release(); L.hdist.release(); L.hrows.release();
    L.hbitmaps.release(); L.hfilters.release(); L.hin.release(); L.hout.release();
    L.ethr.release(); L.seed_keys.release(); L.ecount.release(); L.elist.release();
    L.qn.release(); L.qnorm2.release(); L.cand.release(); L.fail_list.release(); L.counters.release();
    L.extra_cnt.release(); L.timing.release(); L.pin.release();
    for (cudaEvent_t e : {L.ev_in, L.ev_done, L.ev0, L.ev1, L.ev_scored, L.ev_post})
      if (e) cudaEventDestroy(e);
    if (L.pstream) {
      cudaStreamSynchronize(L.pstream);
      cudaStreamDestroy(L.pstream);
    }
    if (L.stream) cudaStreamDestroy(L.stream);
  }
  for (void* p : xopened) cudaIpcCloseMemHandle(p);
  if (xbuf) cudaFree(xbuf);
  xwords.release(); xerr_pin.release();
  w_sel_dist.release(); w_exact.release(); w_maxdist.release();
  w_out_dist.release(); w_pair_sim.release(); w_sel_rows.release();
  w_extra_items.release(); w_qgroup.release(); w_out_count.release();
  w_idx.release(); w_sel_bound.release(); w_queries.release(); w_stage.release(); w_bitmaps.release();
  w_live.release(); w_member.release(); w_out_rows.release(); w_filters.release(); w_pairs.release();
  w_pair_count.release(); w_gather.release();
  w_bad.release(); w_dst.release(); w_cidx.release();
  w_out_pairs.release(); w_jcounts.release(); w_joffsets.release(); w_run_cut.release(); w_jtotal.release();
  w_kind_c.release(); w_jbits.release(); w_ghist.release(); w_keys.release(); w_jstate.release(); w_out_sim.release();
  jpin.release();
  for (cudaEvent_t e : {jev0, jev1, jchunk[0], jchunk[1]})
    if (e) cudaEventDestroy(e);
  if (jstream) cudaStreamDestroy(jstream);
  pin_in.release(); pin_out.release();
  if (stream) cudaStreamDestroy(stream);
}

// Copies host metadata mirrors [row0, row0+count) to the device.
int omr_store::push_meta(int64_t row0, int64_t count) {
  if (count <= 0) return 0;
  const size_t c = (size_t)count;
  OMR_CUDA_OK(cudaMemcpyAsync(d_ws + row0, h_ws.data() + row0, c * 4, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_user + row0, h_user.data() + row0, c * 4, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_agent + row0, h_agent.data() + row0, c * 4, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_kind + row0, h_kind.data() + row0, c * 2, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_purpose + row0, h_purpose.data() + row0, c * 2, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_conf + row0, h_conf.data() + row0, c * 4, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_flags + row0, h_flags.data() + row0, c, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(d_valid + row0, h_valid.data() + row0, c * 8, cudaMemcpyHostToDevice, stream));
  // the row's entity uuid as two big-endian words, so that an unsigned compare on the device is Postgres' memcmp
  std::vector<unsigned long long> ek(2 * c);
  for (size_t i = 0; i < c; i++) {
    uint8_t b[16];
    key_to(entity_id[h_entity[row0 + (int64_t)i]], b);
    unsigned long long hi = 0, lo = 0;
    for (int j = 0; j < 8; j++) {
      hi = (hi << 8) | b[j];
      lo = (lo << 8) | b[8 + j];
    }
    ek[2 * i] = hi;
    ek[2 * i + 1] = lo;
  }
  OMR_CUDA_OK(cudaMemcpyAsync(d_entity_key + 2 * row0, ek.data(), c * 16, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaStreamSynchronize(stream));
  live_dirty = true;
  return 0;
}

int omr_store::push_flag(int64_t row) {
  OMR_CUDA_OK(cudaMemcpy(d_flags + row, &h_flags[row], 1, cudaMemcpyHostToDevice));
  live_dirty = true;
  return 0;
}

// Streams `count` host vectors through a bounded staging buffer into rows [row0, row0+count), or into the rows
// `dst` names (host array, any order).  pgvector refuses NaN and infinite elements when a vector is parsed; the
// ingest kernel flags them and the call fails with pgvector's message (the rows written so far are the caller's to
// discard: bulk loads write beyond `n` before they publish anything).
int omr_store::append(int64_t count, const float* vecs, int64_t row0, const int64_t* dst) {
  const int64_t chunk = std::max<int64_t>(1, (int64_t)(64u << 20) / ((int64_t)dim * 4));
  int rc = w_stage.reserve((size_t)std::min(chunk, count) * dim);
  if (rc) return rc;
  if ((rc = w_bad.reserve(1))) return rc;
  if (dst && (rc = w_dst.reserve((size_t)std::min(chunk, count)))) return rc;
  OMR_CUDA_OK(cudaMemsetAsync(w_bad.p, 0, 4, stream));
  for (int64_t off = 0; off < count; off += chunk) {
    const int64_t c = std::min(chunk, count - off);
    OMR_CUDA_OK(cudaMemcpyAsync(w_stage.p, vecs + off * dim, (size_t)c * dim * 4, cudaMemcpyHostToDevice, stream));
    if (dst) OMR_CUDA_OK(cudaMemcpyAsync(w_dst.p, dst + off, (size_t)c * 8, cudaMemcpyHostToDevice, stream));
    rc = launch_ingest_rows(w_stage.p, c, dim, pitch, row0 + off, dst ? w_dst.p : nullptr, d_bf16, d_f32, d_norm2, w_bad.p,
                            stream);
    if (rc) return rc;
    st_launches++;
    OMR_CUDA_OK(cudaStreamSynchronize(stream));
  }
  int32_t bad = 0;
  OMR_CUDA_OK(cudaMemcpy(&bad, w_bad.p, 4, cudaMemcpyDeviceToHost));
  if (bad & 1) return fail_msg(OMR_E_INVALID, "memory: NaN not allowed in vector");
  if (bad & 2) return fail_msg(OMR_E_INVALID, "memory: infinite value not allowed in vector");
  return 0;
}

// A hard-deleted row goes on the free list (once); omr_upsert hands the slot to the next new observation.
void omr_store::release_row(int64_t r) {
  if (h_flags[r] & kRowFree) return;
  h_flags[r] |= kRowFree;
  free_rows.push_back(r);
}

// Takes a slot off the free list for a NEW observation: the old tenant leaves the identity maps.  -1 = none free.
int64_t omr_store::recycle_row(std::vector<int64_t>* touched) {
  while (!free_rows.empty()) {
    const int64_t r = free_rows.back();
    free_rows.pop_back();
    if (r >= n || !(h_flags[r] & kRowFree)) continue;   // resurrected in place by its own observation id, or compacted away
    auto it = obs_row.find(obs_id[r]);
    if (it != obs_row.end() && it->second == r) obs_row.erase(it);
    const uint32_t ent = h_entity[r];
    auto& rows = entity_rows[ent];
    rows.erase(std::remove(rows.begin(), rows.end(), r), rows.end());
    if (entity_latest[ent] == r) {
      // the entity's representative is its newest live embedded row, as after a merge
      int64_t rep = -1;
      for (int64_t x : rows)
        if (!(h_flags[x] & kRowDeleted) && x > rep) rep = x;
      entity_latest[ent] = rep;
      if (rep >= 0) {
        h_flags[rep] |= kRowEntityRep;
        touched->push_back(rep);
      }
    }
    h_flags[r] = 0;
    return r;
  }
  return -1;
}


// Compaction: every hard-deleted row leaves the shard, the rows behind it move down (order kept), and the
// identity maps follow.  Rows only ever move to a lower index, so the matrices are rewritten in place, a
// chunk at a time through a staging buffer: the sources of a later chunk all lie behind the chunk being written.
int omr_store::compact(int64_t* reclaimed) {
  std::vector<int32_t> keep;
  keep.reserve((size_t)n);
  for (int64_t r = 0; r < n; r++)
    if (!(h_flags[r] & kRowDeleted)) keep.push_back((int32_t)r);
  const int64_t m = (int64_t)keep.size();
  *reclaimed = n - m;
  free_rows.clear();
  if (m == n) return 0;
  int64_t f = 0;
  while (f < m && keep[(size_t)f] == f) f++;   // rows before the first hole stay where they are
  int rc;
  if (m > f) {
    if ((rc = w_cidx.reserve((size_t)(m - f)))) return rc;
    OMR_CUDA_OK(cudaMemcpyAsync(w_cidx.p, keep.data() + f, (size_t)(m - f) * 4, cudaMemcpyHostToDevice, stream));
    const int64_t chunk = std::max<int64_t>(256, (int64_t)(64u << 20) / ((int64_t)pitch * 4));
    if ((rc = w_stage.reserve((size_t)std::min(chunk, m - f) * pitch))) return rc;
    struct Col { void* base; size_t row_bytes; };
    const Col cols[] = {{d_f32, (size_t)pitch * 4}, {d_bf16, (size_t)pitch * 2}, {d_norm2, 8}};
    for (const Col& c : cols) {
      for (int64_t d0 = f; d0 < m; d0 += chunk) {
        const int64_t cnt = std::min(chunk, m - d0);
        if ((rc = launch_gather_row_bytes(c.base, c.row_bytes, w_cidx.p + (d0 - f), cnt, w_stage.p, stream))) return rc;
        OMR_CUDA_OK(cudaMemcpyAsync(reinterpret_cast<uint8_t*>(c.base) + (size_t)d0 * c.row_bytes, w_stage.p,
                                    (size_t)cnt * c.row_bytes, cudaMemcpyDeviceToDevice, stream));
        st_launches++;
      }
    }
  }
  // the vacated tail reads as empty rows
  OMR_CUDA_OK(cudaMemsetAsync(d_flags + m, 0, (size_t)(n - m), stream));
  OMR_CUDA_OK(cudaMemsetAsync(d_norm2 + m, 0, (size_t)(n - m) * 8, stream));
  OMR_CUDA_OK(cudaMemsetAsync(d_bf16 + (size_t)m * pitch, 0, (size_t)(n - m) * pitch * 2, stream));
  OMR_CUDA_OK(cudaStreamSynchronize(stream));
  // host mirrors and identity maps
  std::vector<int64_t> newpos((size_t)n, -1);
  for (int64_t i = 0; i < m; i++) newpos[(size_t)keep[(size_t)i]] = i;
  for (int64_t i = f; i < m; i++) {
    const int64_t o = keep[(size_t)i];
    h_ws[i] = h_ws[o]; h_user[i] = h_user[o]; h_agent[i] = h_agent[o]; h_entity[i] = h_entity[o];
    h_kind[i] = h_kind[o]; h_purpose[i] = h_purpose[o]; h_conf[i] = h_conf[o]; h_flags[i] = h_flags[o];
    h_valid[i] = h_valid[o]; obs_id[i] = obs_id[o];
  }
  obs_row.clear();
  for (auto& rows : entity_rows) rows.clear();
  for (int64_t i = 0; i < m; i++) {
    obs_row[obs_id[i]] = i;
    entity_rows[h_entity[i]].push_back(i);
  }
  for (size_t e = 0; e < entity_latest.size(); e++) {
    const int64_t old = entity_latest[e];
    int64_t rep = old >= 0 ? newpos[(size_t)old] : -1;
    if (rep < 0 && !entity_rows[e].empty()) {   // its representative was deleted: the newest surviving row takes over
      rep = entity_rows[e].back();
      h_flags[rep] |= kRowEntityRep;
    }
    entity_latest[e] = rep;
  }
  const size_t t = (size_t)m;
  h_ws.resize(t); h_user.resize(t); h_agent.resize(t); h_entity.resize(t); h_kind.resize(t); h_purpose.resize(t);
  h_conf.resize(t); h_flags.resize(t); h_valid.resize(t); obs_id.resize(t);
  n = m;
  return push_meta(0, n);
}

int omr_store::to_device_filter(const omr_filter& f, DeviceFilter* out, double* maxd) {
  if (!f.workspace_id || !f.workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  DeviceFilter d{};
  if (!ws_names.lookup(f.workspace_id, &d.workspace)) d.workspace = kNoSuchId;
  auto scope = [&](const char* v, int mode, const Interner& in, uint32_t* id, int32_t* m) -> int {
    if (mode < OMR_MATCH_ANY || mode > OMR_MATCH_NULL_OR_EQ) return fail_msg(OMR_E_INVALID, "memory: bad scope match mode %d", mode);
    *m = mode;
    *id = 0;
    if (mode == OMR_MATCH_EQ || mode == OMR_MATCH_NULL_OR_EQ) {
      if (!v) {  // value absent: EQ degrades to "no constraint", NULL_OR_EQ to "IS NULL", as the SQL does
        *m = mode == OMR_MATCH_EQ ? OMR_MATCH_ANY : OMR_MATCH_NULL_ONLY;
      } else if (!in.lookup(v, id)) {
        *id = kNoSuchId;
      }
    }
    return 0;
  };
  int rc = scope(f.virtual_user_id, f.user_mode, user_names, &d.user, &d.user_mode);
  if (rc) return rc;
  rc = scope(f.agent_id, f.agent_mode, agent_names, &d.agent, &d.agent_mode);
  if (rc) return rc;
  if (f.n_kinds > kMaxFilterSet || f.n_purposes > kMaxFilterSet)
    return fail_msg(OMR_E_INVALID, "memory: at most %d kinds / purposes per filter", kMaxFilterSet);
  d.n_kinds = f.n_kinds;
  for (int i = 0; i < f.n_kinds; i++) {
    uint32_t id;
    d.kinds[i] = kind_names.lookup(f.kinds[i], &id) ? (uint16_t)id : (uint16_t)0xFFFF;
  }
  d.n_purposes = f.n_purposes;
  for (int i = 0; i < f.n_purposes; i++) {
    uint32_t id;
    d.purposes[i] = purpose_names.lookup(f.purposes[i], &id) ? (uint16_t)id : (uint16_t)0xFFFF;
  }
  d.min_confidence = f.min_confidence;
  d.now_us = f.now_us;
  *out = d;
  *maxd = (f.max_distance == f.max_distance && f.max_distance < 2.0) ? f.max_distance : INFINITY;
  return 0;
}

// Bitmap of every live row (any workspace, no expiry test): the default when the caller of
// omr_device_search passes no bitmaps.
int omr_store::ensure_live(cudaStream_t st) {
  int rc = w_live.reserve((size_t)words());
  if (rc) return rc;
  if (!live_dirty) return 0;
  rc = w_filters.reserve(64);
  if (rc) return rc;
  DeviceFilter d{};
  d.workspace = kAnyWorkspace;
  d.min_confidence = -INFINITY;
  d.now_us = INT64_MIN;
  OMR_CUDA_OK(cudaMemcpyAsync(w_filters.p + 63, &d, sizeof d, cudaMemcpyHostToDevice, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));  // `d` is a stack object
  rc = launch_build_bitmaps(meta(), n, (int)words(), w_filters.p + 63, 1, w_live.p, st);
  if (rc) return rc;
  st_launches++;
  live_dirty = false;
  return 0;
}

int omr_store::device_search(const float* d_q, int nq, const uint32_t* d_bitmaps, const int32_t* d_qgroup,
                             const double* d_maxdist, int k, int64_t* d_rows, double* d_dist, int32_t* d_count,
                             cudaStream_t st) {
  Lane& L = lanes[0];
  int rc = enqueue(L, d_q, nq, d_bitmaps, d_qgroup, d_maxdist, k, d_rows, d_dist, d_count, st, nullptr);
  if (rc) return rc;
  return complete(L, st);
}

// Waits for every batch of the begin/finish pipeline that is still in flight (before the store mutates).
int omr_store::drain() {
  for (int i = 1; i < kLanes; i++)
    if (lanes[i].busy && !lanes[i].immediate) OMR_CUDA_OK(cudaStreamSynchronize(lanes[i].stream));
  return 0;
}

// Emission-mode scratch of a lane: thresholds, counts, candidate lists and the seed keys.
// Seed = top-2-per-32-row keys of the first row tiles (one wave of the scorer at most, at least enough tiles
// for k keys, a sixteenth of a small store); they give every query its first threshold.
int omr_store::prepare_emit(Lane& L, int qpad, int k, int n_tiles, cudaStream_t st, EmitView* out, int* seed_n,
                            int64_t* seed_kpq) {
  int rc;
  EmitView ev;
  ev.k = k;
  ev.band = 2.0f * margin;
  // entries per query: a 10M-row scan at k = 100 ends with ~4200 on average, and a query that overflows costs an
  // exact scan of the whole shard
  ev.cap = std::min(16384, std::max(2048, pow2ceil_i(96 * k)));
  // measured (profiles/r02_trigger_sweep.txt, r02_refresh_warp_ab.txt): too small a trigger floods the refresh warp
  // on short shards (125k rows, k = 10: scorer 0.39 ms at 32, 0.28 ms at 128), too large a one lets the lists of a
  // large k grow (1.25M x 768, top-100: 570 k queries/s at 256, 590 k at 512, 574 k at 1024)
  ev.trigger = std::max(128, pow2ceil_i(4 * k));
  if (const char* e = getenv("OMR_EMIT_TRIGGER")) ev.trigger = std::max(32, pow2ceil_i(atoi(e)));   // tuning knob
  const int wave = std::max(1, num_sms / std::max(1, qpad / kQueryTile));
  *seed_n = std::min(n_tiles, std::max((k + 7) / 8 + 1, std::min(std::min(20, wave), std::max(1, n_tiles / 16))));
  if (const char* e = getenv("OMR_SEED_TILES"))   // tuning knob; never fewer tiles than hold k keys
    *seed_n = std::min(n_tiles, std::max((k + 7) / 8 + 1, std::min(40, atoi(e))));
  *seed_kpq = (int64_t)*seed_n * (kTileRows / 32) * 2;
  if ((rc = L.ethr.reserve((size_t)qpad))) return rc;
  if ((rc = L.ecount.reserve((size_t)qpad))) return rc;
  const size_t had = L.elist.n;
  if ((rc = L.elist.reserve((size_t)qpad * ev.cap))) return rc;
  if ((rc = L.seed_keys.reserve((size_t)qpad * *seed_kpq))) return rc;
  if (L.elist.n != had || L.elist_dirty)   // fresh memory, or a batch that never reached its post kernel
    OMR_CUDA_OK(cudaMemsetAsync(L.elist.p, 0, L.elist.n * sizeof(unsigned long long), st));
  L.elist_dirty = true;
  ev.thr = L.ethr.p;
  ev.count = L.ecount.p;
  ev.list = L.elist.p;
  *out = ev;
  return 0;
}

// Enqueues one batch on `st` up to (and including) the read-back of its counters; no host wait.
int omr_store::enqueue(Lane& L, const float* d_q, int nq, const uint32_t* d_bitmaps, const int32_t* d_qgroup,
                       const double* d_maxdist, int k, int64_t* d_rows, double* d_dist, int32_t* d_count,
                       cudaStream_t st, bool* fused_out) {
  L.nq = 0;
  L.nfail_known = 0;
  L.timed = false;
  L.phased = false;
  if (fused_out) *fused_out = true;
  if (nq <= 0) return 0;
  if (k <= 0 || k > 4096) return fail_msg(OMR_E_INVALID, "memory: k must be in 1..4096, got %d", k);
  st_searches++;
  st_queries += nq;
  int rc;
  if (n == 0) {
    rc = launch_fill_empty(d_rows, d_dist, d_count, nq, k, st);
    st_launches++;
    return rc;
  }
  if (!d_bitmaps) {
    rc = ensure_live(st);
    if (rc) return rc;
    d_bitmaps = w_live.p;
    d_qgroup = nullptr;
  }
  const int n_tiles = (int)((n + kTileRows - 1) / kTileRows);
  const int wpg = (int)words();
  int K2 = cfg.rescore_width > 0 ? cfg.rescore_width : pow2ceil_i(std::max(32, 2 * k));
  K2 = std::min(std::max(K2, k), 4096);
  const int K2x = K2 + kExtraSlots;
  int S = cfg.sub_bucket > 0 ? cfg.sub_bucket : (k <= 32 ? 128 : (k <= 128 ? 64 : 32));
  // Measured on B200 (profiles/r01_configs_single_gpu.jsonl): the TMA-fed tensor-core scorer streams the
  // matrix at ~6.7 TB/s for any batch up to ~128 queries, above the LDG-based streaming scorer for
  // every batch but a single query.
  // One query: the bulk-copy streaming scorer.  Up to one query tile: the 1-CTA tensor-core scorer (HBM-bound
  // there).  More: CTA pairs, whose lower L2->SM traffic per flop holds higher clocks under the power cap
  // (profiles/r02_*: 466 k vs 439 k queries/s at 1M x 1536, 400 k vs 381 k sustained).
  int path = cfg.force_path > 0 ? cfg.force_path : (nq <= 1 ? 1 : (nq <= kQueryTile ? 2 : 4));
  // Expected rescued rows per query ~ m^2 S^2 / (2n), m = rows whose approximate score clears the
  // certificate's threshold (the k answers plus those inside the margin; 2k allows for as many again).
  // When even the smallest sub-bucket cannot keep that small (large k on a small store), the two-stage
  // scheme has nothing to offer: scan exactly.
  if (cfg.sub_bucket == 0 && path != 3) {
    const double m2 = 4.0 * (double)k * k;
    while (S > 32 && m2 * S * S / (2.0 * (double)n) > 256.0) S >>= 1;
    if (cfg.force_path == 0 && m2 * S * S / (2.0 * (double)n) > 1024.0) path = 3;
  }
  const int qtile = path == 4 ? 2 * kQueryTile : kQueryTile;
  const int qpad = (path == 2 || path == 4) ? ((nq + qtile - 1) / qtile) * qtile : std::max(nq, 8);
  const int64_t kpq = (int64_t)n_tiles * (kTileRows / S) * 2;
  last_path = path; last_S = S; last_K2 = K2;

  // Tensor-core paths with k <= 256 run in emission mode: no key array at all, the scorer appends the rows that
  // clear a per-query running threshold to a short list (k_score_umma.cu) and one kernel finishes from it.
  const bool emit = (path == 2 || path == 4) && !legacy_post && emit_supported(k);
  const bool fused = path != 3 && !legacy_post && (emit || post_fused_supported(K2, k));
  if (fused_out) *fused_out = fused;
  EmitView ev;
  int seed_n = 0;
  int64_t seed_kpq = 0;
  if (emit && (rc = prepare_emit(L, qpad, k, n_tiles, st, &ev, &seed_n, &seed_kpq))) return rc;
  if ((rc = L.qn.reserve((size_t)qpad * pitch))) return rc;
  if ((rc = L.qnorm2.reserve((size_t)qpad))) return rc;
  if ((rc = L.fail_list.reserve((size_t)qpad))) return rc;
  if ((rc = L.extra_cnt.reserve((size_t)qpad))) return rc;
  if (!fused) {
    if ((rc = w_sel_rows.reserve((size_t)qpad * K2x))) return rc;
    if ((rc = w_sel_dist.reserve((size_t)qpad * K2x))) return rc;
    if ((rc = w_sel_bound.reserve((size_t)qpad))) return rc;
    if ((rc = w_extra_items.reserve((size_t)qpad * kExtraSlots))) return rc;
  }
  if (path != 3 && !emit && (rc = L.cand.reserve((size_t)qpad * kpq))) return rc;

  OMR_CUDA_OK(cudaMemcpyAsync(L.counters.p, counters_init.p, 16 * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  OMR_CUDA_OK(cudaMemsetAsync(L.extra_cnt.p, 0, (size_t)qpad * sizeof(int32_t), st));
  int32_t* extra_total = L.counters.p;
  int32_t* fail_count = L.counters.p + 1;
  const ShardView sh = view();

  rc = launch_prep_queries(d_q, nq, qpad, dim, pitch, L.qn.p, L.qnorm2.p, L.counters.p + 2, st);
  if (rc) return rc;
  st_launches++;

  FinalizeArgs fa;
  fa.nq = nq; fa.K2 = K2; fa.K2x = K2x; fa.k = k; fa.S = S;
  fa.sel_rows = w_sel_rows.p; fa.sel_dist = w_sel_dist.p; fa.max_dist = d_maxdist; fa.sel_bound = w_sel_bound.p;
  fa.cand_keys = L.cand.p; fa.keys_per_query = kpq; fa.margin = margin;
  fa.bitmaps = d_bitmaps; fa.words_per_group = wpg; fa.query_group = d_qgroup;
  fa.row_base = cfg.row_base; fa.row_stride = cfg.row_stride;
  fa.out_rows = d_rows; fa.out_dist = d_dist; fa.out_count = d_count;
  fa.extra_cnt = L.extra_cnt.p; fa.extra_items = w_extra_items.p; fa.extra_total = extra_total;
  fa.fail_list = L.fail_list.p; fa.fail_count = fail_count;

  L.d_q = d_q; L.d_bitmaps = d_bitmaps; L.d_qgroup = d_qgroup; L.nq = nq; L.k = k; L.K2 = K2; L.K2x = K2x; L.qpad = qpad;
  L.fa = fa;
  if (path == 3) {
    rc = launch_iota(L.fail_list.p, nq, st);
    if (rc) return rc;
    st_launches++;
    L.nfail_known = nq;
    OMR_CUDA_OK(cudaMemcpyAsync(L.pin.p, L.counters.p, 3 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    return 0;
  }
  L.nfail_known = -1;
  {
    if (profiling && !emit) OMR_CUDA_OK(cudaEventRecord(L.ev0, st));
    if (path == 1) {
      // the streaming scorer takes 8 queries per pass over the matrix
      for (int q0 = 0; q0 < nq; q0 += 8) {
        const int c = std::min(8, nq - q0);
        rc = launch_score_stream(sh, L.qn.p + (size_t)q0 * pitch, c, n_tiles, S, d_bitmaps, wpg,
                                 d_qgroup ? d_qgroup + q0 : nullptr, L.cand.p + (size_t)q0 * kpq, num_sms, st);
        if (rc) return rc;
        st_launches++;
      }
    } else {
      if (L.tmap_q_ptr != L.qn.p || L.tmap_q_rows != L.qn.n / pitch) {
        rc = make_tmap(&L.tmap_q, L.qn.p, L.qn.n / pitch, kQueryTile);
        if (rc) return rc;
        L.tmap_q_ptr = L.qn.p;
        L.tmap_q_rows = L.qn.n / pitch;
      }
      if (emit) {
        // The three kernels are chained by programmatic dependent launch: the full-store scorer is scheduled onto
        // SMs as they come free and runs its loads and MMAs while the seed finishes; only its threshold reads wait.
        // Nothing may be enqueued between them, so the scorer is timed by its own %globaltimer stamps (the six words
        // behind the lane's counters, initialised and read back with them): profiling adds no stream operation.
        const bool pdl = use_pdl;
        unsigned long long* timer = profiling ? reinterpret_cast<unsigned long long*>(L.counters.p + 4) : nullptr;
        rc = launch_score_umma(L.tmap_q, tmap_rows_b, nq, qpad / kQueryTile, seed_n, pitch / kBlockK, 32, d_bitmaps, wpg,
                               d_qgroup, L.seed_keys.p, num_sms, st);
        if (rc) return rc;
        if ((rc = launch_emit_seed(L.seed_keys.p, seed_kpq, L.qnorm2.p, nq, qpad, ev, pdl, st))) return rc;
        st_launches += 2;
        if (path == 4)
          rc = launch_score_umma2_emit(L.tmap_q, tmap_rows_a, nq, qpad / (2 * kQueryTile), n_tiles, pitch / kBlockK,
                                       d_bitmaps, wpg, d_qgroup, ev, timer, num_sms, pdl, st);
        else
          rc = launch_score_umma_emit(L.tmap_q, tmap_rows_b, nq, qpad / kQueryTile, n_tiles, pitch / kBlockK, d_bitmaps,
                                      wpg, d_qgroup, ev, timer, num_sms, pdl, st);
        L.phased = timer != nullptr;
      } else if (path == 4)
        rc = launch_score_umma2(L.tmap_q, tmap_rows_a, nq, qpad / (2 * kQueryTile), n_tiles, pitch / kBlockK, S, d_bitmaps, wpg,
                                d_qgroup, L.cand.p, num_sms, st);
      else
        rc = launch_score_umma(L.tmap_q, tmap_rows_b, nq, qpad / kQueryTile, n_tiles, pitch / kBlockK, S, d_bitmaps, wpg,
                               d_qgroup, L.cand.p, num_sms, st);
      if (rc) return rc;
      st_launches++;
    }
    if (profiling && !emit) OMR_CUDA_OK(cudaEventRecord(L.ev1, st));
    if (emit) {
      EmitPostArgs pa;
      pa.sh = sh; pa.q = d_q; pa.qnorm2 = L.qnorm2.p; pa.ev = ev;
      pa.cand = emit_post_capacity(pitch);
      pa.nq = nq; pa.k = k; pa.K2 = cfg.rescore_width > 0 ? std::min(std::max(K2, k), pa.cand) : std::max(16, k + 6);
      pa.max_dist = d_maxdist; pa.margin = margin;
      pa.row_base = cfg.row_base; pa.row_stride = cfg.row_stride;
      pa.out_rows = d_rows; pa.out_dist = d_dist; pa.out_count = d_count;
      pa.fail_list = L.fail_list.p; pa.fail_count = fail_count; pa.rescued_total = extra_total;
      pa.listed_total = L.counters.p + 3;
      if (L.phased) pa.timer = reinterpret_cast<unsigned long long*>(L.counters.p + 4) + 4;
      OMR_CUDA_OK(cudaEventRecord(L.ev_scored, st));
      OMR_CUDA_OK(cudaStreamWaitEvent(L.pstream, L.ev_scored, 0));
      if ((rc = launch_emit_post(pa, L.pstream))) return rc;
      OMR_CUDA_OK(cudaEventRecord(L.ev_post, L.pstream));
      OMR_CUDA_OK(cudaStreamWaitEvent(st, L.ev_post, 0));
      st_launches += 1;
      L.elist_dirty = false;
    } else if (fused) {
      PostArgs pa;
      pa.sh = sh; pa.q = d_q; pa.qnorm2 = L.qnorm2.p; pa.cand_keys = L.cand.p; pa.keys_per_query = kpq;
      // the fused kernel tops its candidate list up adaptively, so its first pass can be narrow
      pa.nq = nq; pa.S = S; pa.K2 = cfg.rescore_width > 0 ? K2 : std::min(K2, std::max(16, k + 6)); pa.k = k; pa.max_dist = d_maxdist; pa.margin = margin;
      pa.bitmaps = d_bitmaps; pa.words_per_group = wpg; pa.query_group = d_qgroup;
      pa.row_base = cfg.row_base; pa.row_stride = cfg.row_stride;
      pa.out_rows = d_rows; pa.out_dist = d_dist; pa.out_count = d_count;
      pa.fail_list = L.fail_list.p; pa.fail_count = fail_count; pa.rescued_total = extra_total;
      // small enough that a block (fp64 query + sort arrays + 4 KB) stays under ~30 KB and fits beside a
      // scorer CTA of the next batch in flight
      pa.sort_cap = k <= 16 ? 512 : (k <= 128 ? 1024 : 2048);
      if (post_timing) {
        if ((rc = L.timing.reserve((size_t)qpad * 8))) return rc;
        pa.timing = L.timing.p;
      }
      if ((rc = launch_post(pa, st))) return rc;
      st_launches += 1;
    } else {
      if ((rc = launch_select_keys(L.cand.p, nq, kpq, S, K2, K2x, w_sel_rows.p, w_sel_bound.p, st))) return rc;
      if ((rc = launch_rescore(sh, d_q, L.qnorm2.p, nq, K2, K2x, w_sel_rows.p, w_sel_dist.p, st))) return rc;
      if ((rc = launch_finalize(fa, st))) return rc;
      if ((rc = launch_rescore_items(sh, d_q, L.qnorm2.p, K2x, w_extra_items.p, extra_total, w_sel_rows.p, w_sel_dist.p,
                                     num_sms, st)))
        return rc;
      if ((rc = launch_finalize_extra(fa, st))) return rc;
      st_launches += 5;
    }
    OMR_CUDA_OK(cudaMemcpyAsync(L.pin.p, L.counters.p, (L.phased ? 16 : 4) * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    L.timed = profiling && !emit;
  }
  return 0;
}

// Host side of a batch: waits for `st`, reads the counters and runs the exact fallback for the
// (rare) queries whose certificate failed.
int omr_store::complete(Lane& L, cudaStream_t st) {
  L.nfail_seen = 0;
  if (L.nq <= 0) return 0;
  int rc;
  int nfail = L.nfail_known;
  L.nfail_seen = 0;
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  // pgvector refuses non-finite elements when the query vector is parsed, so the reference's statement
  // fails as a whole; the query-preparation kernel flags them
  if (L.pin.p[2] & 1) return fail_msg(OMR_E_INVALID, "memory: NaN not allowed in vector");
  if (L.pin.p[2] & 2) return fail_msg(OMR_E_INVALID, "memory: infinite value not allowed in vector");
  if (nfail < 0) {
    st_rescued += L.pin.p[0];
    st_listed += L.pin.p[3];
    nfail = L.pin.p[1];
    if (L.timed) {   // key-table paths: stream events around the scorer
      float ms = 0.0f;
      OMR_CUDA_OK(cudaEventElapsedTime(&ms, L.ev0, L.ev1));
      scorer_ms_sum += ms;
      scorer_timed++;
    }
    if (L.phased) {   // emission mode: the kernels' own %globaltimer stamps, read back with the counters
      unsigned long long t01[4];
      memcpy(t01, L.pin.p + 4, 32);
      if (t01[1] > t01[0]) {
        kernel_ns_sum += (double)(t01[1] - t01[0]);
        kernel_start_skew_ns += (double)(t01[2] - t01[0]);
        kernel_end_skew_ns += (double)(t01[1] - t01[3]);
        kernel_ns_n++;
        // idle between consecutive scorers on this GPU (batches are finished in the order they began)
        if (last_main_end_ns && t01[0] > last_main_end_ns && t01[0] - last_main_end_ns < 50000000ull)
          gap_ns_sum += (double)(t01[0] - last_main_end_ns);
        last_main_end_ns = t01[1];
        unsigned long long tp[2];
        memcpy(tp, L.pin.p + 12, 16);
        if (tp[1] > tp[0]) {
          post_ns_sum += (double)(tp[1] - tp[0]);
          post_lag_ns_sum += (double)(tp[1] > t01[1] ? tp[1] - t01[1] : 0);
        }
      }
    }
  }
  L.nfail_seen = nfail;
  if (nfail > 0) {
    // Exact fallback: score every eligible row of the failing queries in fp64.
    const int nq = L.nq, k = L.k, K2 = L.K2, K2x = L.K2x, qpad = L.qpad;
    const float* d_q = L.d_q;
    const uint32_t* d_bitmaps = L.d_bitmaps;
    const int32_t* d_qgroup = L.d_qgroup;
    const int wpg = (int)words();
    const ShardView sh = view();
    FinalizeArgs fa = L.fa;
    (void)nq;
    st_fallback += nfail;
    if ((rc = w_sel_rows.reserve((size_t)qpad * K2x))) return rc;   // the fallback's tail uses the candidate-slot layout
    if ((rc = w_sel_dist.reserve((size_t)qpad * K2x))) return rc;
    fa.sel_rows = w_sel_rows.p;
    fa.sel_dist = w_sel_dist.p;
    const int64_t n_pad = (int64_t)wpg * 32;
    const int slots = (int)std::max<int64_t>(1, std::min<int64_t>(nfail, (int64_t)(256u << 20) / (n_pad * 8)));
    if ((rc = w_exact.reserve((size_t)slots * n_pad))) return rc;
    FinalizeArgs fx = fa;
    fx.cand_keys = nullptr;
    for (int s0 = 0; s0 < nfail; s0 += slots) {
      const int c = std::min(slots, nfail - s0);
      const int32_t* list = L.fail_list.p + s0;
      if ((rc = launch_exact_scan(sh, d_q, L.qnorm2.p, list, c, d_bitmaps, wpg, d_qgroup, w_exact.p, st))) return rc;
      if ((rc = launch_select_exact(w_exact.p, n_pad, list, c, K2, K2x, k, w_sel_rows.p, w_sel_dist.p, st))) return rc;
      fx.qlist = list;
      fx.nlist = c;
      if ((rc = launch_finalize(fx, st))) return rc;
      st_launches += 3;
    }
    OMR_CUDA_OK(cudaStreamSynchronize(st));
  }
  return 0;
}

void omr_store::fill_hit(omr_hit* h, int64_t grow, double dist) const {
  const int64_t stride = cfg.row_stride > 0 ? cfg.row_stride : 1;
  const int64_t local = (grow - cfg.row_base) / stride;
  key_to(obs_id[local], h->observation_id);
  key_to(entity_id[h_entity[local]], h->entity_id);
  h->distance = dist;
  h->row = grow;
}

int omr_store::host_search(const float* queries, int nq, const omr_filter* filters, int n_filters,
                           const int32_t* query_filter, int k, omr_hit* hits, int32_t* counts, int64_t* soa_rows,
                           double* soa_dist) {
  if (nq <= 0) return 0;
  if (n_filters <= 0 || !filters) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  int rc;
  std::vector<DeviceFilter> df((size_t)n_filters);
  std::vector<double> fmax((size_t)n_filters);
  for (int i = 0; i < n_filters; i++)
    if ((rc = to_device_filter(filters[i], &df[i], &fmax[i]))) return rc;
  // staging layout in pinned memory: queries | query_group | max_dist
  const size_t qbytes = (size_t)nq * dim * 4;
  const size_t in_bytes = qbytes + (size_t)nq * 4 + (size_t)nq * 8 + 16;
  if ((rc = pin_in.reserve(in_bytes))) return rc;
  float* pq = reinterpret_cast<float*>(pin_in.p);
  double* pm = reinterpret_cast<double*>(pin_in.p + ((qbytes + 7) & ~size_t(7)));
  int32_t* pg = reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(pm) + (size_t)nq * 8);
  // A caller whose query buffer is already page-locked (cudaHostAlloc / cudaHostRegister) is copied from
  // directly; pageable memory goes through the store's pinned staging buffer.
  const float* qsrc = pq;
  {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, queries) == cudaSuccess && attr.type == cudaMemoryTypeHost) qsrc = queries;
    else { cudaGetLastError(); memcpy(pq, queries, qbytes); }
  }
  for (int i = 0; i < nq; i++) {
    const int g = query_filter ? query_filter[i] : 0;
    if (g < 0 || g >= n_filters) return fail_msg(OMR_E_INVALID, "memory: query %d names filter %d of %d", i, g, n_filters);
    pg[i] = g;
    pm[i] = fmax[g];
  }
  if ((rc = w_queries.reserve((size_t)nq * dim))) return rc;
  if ((rc = w_qgroup.reserve((size_t)nq))) return rc;
  if ((rc = w_maxdist.reserve((size_t)nq))) return rc;
  if ((rc = w_filters.reserve((size_t)std::max(64, n_filters + 1)))) return rc;
  if ((rc = w_bitmaps.reserve((size_t)n_filters * words()))) return rc;
  if ((rc = w_out_rows.reserve((size_t)nq * k))) return rc;
  if ((rc = w_out_dist.reserve((size_t)nq * k))) return rc;
  if ((rc = w_out_count.reserve((size_t)nq))) return rc;
  OMR_CUDA_OK(cudaMemcpyAsync(w_queries.p, qsrc, qbytes, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(w_qgroup.p, pg, (size_t)nq * 4, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(w_maxdist.p, pm, (size_t)nq * 8, cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(w_filters.p, df.data(), (size_t)n_filters * sizeof(DeviceFilter), cudaMemcpyHostToDevice, stream));
  OMR_CUDA_OK(cudaStreamSynchronize(stream));  // df is pageable
  if ((rc = launch_build_bitmaps(meta(), n, (int)words(), w_filters.p, n_filters, w_bitmaps.p, stream))) return rc;
  st_launches++;
  if ((rc = device_search(w_queries.p, nq, w_bitmaps.p, w_qgroup.p, w_maxdist.p, k, w_out_rows.p, w_out_dist.p,
                          w_out_count.p, stream)))
    return rc;
  const size_t ob = (size_t)nq * k * 16 + (size_t)nq * 4;
  if ((rc = pin_out.reserve(ob))) return rc;
  int64_t* orow = reinterpret_cast<int64_t*>(pin_out.p);
  double* odist = reinterpret_cast<double*>(pin_out.p + (size_t)nq * k * 8);
  int32_t* ocnt = reinterpret_cast<int32_t*>(pin_out.p + (size_t)nq * k * 16);
  OMR_CUDA_OK(cudaMemcpyAsync(orow, w_out_rows.p, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(odist, w_out_dist.p, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, stream));
  OMR_CUDA_OK(cudaMemcpyAsync(ocnt, w_out_count.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, stream));
  OMR_CUDA_OK(cudaStreamSynchronize(stream));
  if (soa_rows) {
    memcpy(soa_rows, orow, (size_t)nq * k * 8);
    memcpy(soa_dist, odist, (size_t)nq * k * 8);
    memcpy(counts, ocnt, (size_t)nq * 4);
    return 0;
  }
  for (int i = 0; i < nq; i++) {
    counts[i] = ocnt[i];
    for (int j = 0; j < ocnt[i]; j++) fill_hit(&hits[(size_t)i * k + j], orow[(size_t)i * k + j], odist[(size_t)i * k + j]);
  }
  return 0;
}

// pgvector refuses non-finite elements when a vector is parsed ("NaN not allowed in vector", "infinite value
// not allowed in vector").  Single vectors are checked here on the host; query batches on the device.
static int check_finite(const float* v, size_t count) {
  for (size_t i = 0; i < count; i++) {
    if (v[i] != v[i]) return fail_msg(OMR_E_INVALID, "memory: NaN not allowed in vector");
    if (v[i] > 3.402823466e38f || v[i] < -3.402823466e38f)
      return fail_msg(OMR_E_INVALID, "memory: infinite value not allowed in vector");
  }
  return 0;
}

static bool same_cstr(const char* a, const char* b) { return (!a && !b) || (a && b && strcmp(a, b) == 0); }
static bool same_filter(const omr_filter& a, const omr_filter& b) {
  if (!same_cstr(a.workspace_id, b.workspace_id) || !same_cstr(a.virtual_user_id, b.virtual_user_id) ||
      !same_cstr(a.agent_id, b.agent_id) || a.user_mode != b.user_mode || a.agent_mode != b.agent_mode ||
      a.n_kinds != b.n_kinds || a.n_purposes != b.n_purposes || a.min_confidence != b.min_confidence ||
      a.now_us != b.now_us)
    return false;
  if (!(a.max_distance == b.max_distance || (a.max_distance != a.max_distance && b.max_distance != b.max_distance)))
    return false;
  for (int i = 0; i < a.n_kinds; i++)
    if (!same_cstr(a.kinds[i], b.kinds[i])) return false;
  for (int i = 0; i < a.n_purposes; i++)
    if (!same_cstr(a.purposes[i], b.purposes[i])) return false;
  return true;
}

// Runs the requests of one batch as a single search (their callers are blocked in submit(), so every
// pointer they passed stays valid).  Identical filters share one bitmap.
void omr_store::run_batch(std::vector<HostReq*>& batch) {
  std::lock_guard<std::mutex> lk(mu);
  cudaSetDevice(cfg.device);
  st_host_batches++;
  auto run_one = [&](HostReq* r) {
    r->rc = host_search(r->q, r->nq, r->filters, r->n_filters, r->qf, r->k, r->hits, r->counts, r->rows, r->dist);
    if (r->rc) r->err = g_err;
  };
  if (batch.size() == 1) {
    run_one(batch[0]);
    return;
  }
  const int k = batch[0]->k;
  int tq = 0;
  for (HostReq* r : batch) tq += r->nq;
  std::vector<float> q((size_t)tq * dim);
  std::vector<omr_filter> filters;
  std::vector<int32_t> qf((size_t)tq);
  int off = 0;
  bool ok = true;
  for (HostReq* r : batch) {
    memcpy(q.data() + (size_t)off * dim, r->q, (size_t)r->nq * dim * 4);
    std::vector<int> map((size_t)std::max(r->n_filters, 0));
    for (int f = 0; f < r->n_filters; f++) {
      int hit = -1;
      for (size_t g = 0; g < filters.size() && hit < 0; g++)
        if (same_filter(filters[g], r->filters[f])) hit = (int)g;
      if (hit < 0) {
        hit = (int)filters.size();
        filters.push_back(r->filters[f]);
      }
      map[f] = hit;
    }
    for (int i = 0; i < r->nq; i++) {
      const int f = r->qf ? r->qf[i] : 0;
      if (f < 0 || f >= r->n_filters) { ok = false; break; }
      qf[(size_t)off + i] = map[f];
    }
    off += r->nq;
  }
  std::vector<int64_t> rows((size_t)tq * k);
  std::vector<double> dist((size_t)tq * k);
  std::vector<int32_t> cnt((size_t)tq);
  int rc = ok ? host_search(q.data(), tq, filters.data(), (int)filters.size(), qf.data(), k, nullptr, cnt.data(),
                            rows.data(), dist.data())
              : OMR_E_INVALID;
  if (rc) {  // one bad request must not fail its neighbours: run them one by one for exact attribution
    for (HostReq* r : batch) run_one(r);
    return;
  }
  off = 0;
  for (HostReq* r : batch) {
    for (int i = 0; i < r->nq; i++) {
      const size_t src = (size_t)(off + i) * k, dst = (size_t)i * k;
      r->counts[i] = cnt[(size_t)off + i];
      if (r->rows) {
        memcpy(r->rows + dst, rows.data() + src, (size_t)k * 8);
        memcpy(r->dist + dst, dist.data() + src, (size_t)k * 8);
      } else {
        for (int j = 0; j < r->counts[i]; j++) fill_hit(&r->hits[dst + j], rows[src + j], dist[src + j]);
      }
    }
    r->rc = 0;
    off += r->nq;
  }
}

int omr_store::submit(HostReq& req) {
  st_host_calls++;
  std::unique_lock<std::mutex> lk(bq_mu);
  bq.push_back(&req);
  while (!req.done) {
    if (!bq_leader) {
      bq_leader = true;
      std::vector<HostReq*> batch;
      const int k = bq.front()->k;
      int tq = 0, tf = 0;
      for (auto it = bq.begin(); it != bq.end();) {
        HostReq* r = *it;
        if (r->k == k && (batch.empty() || (tq + r->nq <= 4096 && tf + r->n_filters <= 64))) {
          batch.push_back(r);
          tq += r->nq;
          tf += r->n_filters;
          it = bq.erase(it);
        } else {
          ++it;
        }
      }
      lk.unlock();
      run_batch(batch);
      lk.lock();
      for (HostReq* r : batch) r->done = true;
      bq_leader = false;
      bq_cv.notify_all();
    } else {
      bq_cv.wait(lk);
    }
  }
  if (req.rc) g_err = req.err;
  return req.rc;
}

// ================================ C ABI =========================================

extern "C" {

const char* omr_last_error(void) { return g_err.c_str(); }

int omr_open(const omr_config* cfg, omr_store** out) {
  if (!cfg || !out) return fail_msg(OMR_E_INVALID, "memory: omr_open: null argument");
  *out = nullptr;
  if (cfg->dim < 1 || cfg->dim > OMR_MAX_DIM)
    return fail_msg(OMR_E_DIM, "memory: embedding dimension %d outside 1..%d", cfg->dim, OMR_MAX_DIM);
  if (cfg->capacity < 1) return fail_msg(OMR_E_INVALID, "memory: capacity must be positive");
  if (cfg->sub_bucket != 0 && cfg->sub_bucket != 32 && cfg->sub_bucket != 64 && cfg->sub_bucket != 128 &&
      cfg->sub_bucket != 256)
    return fail_msg(OMR_E_INVALID, "memory: sub_bucket must be 0, 32, 64, 128 or 256");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail_msg(OMR_E_CUDA, "memory: gpu: no CUDA device (%s); this engine has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (cfg->device < 0 || cfg->device >= ndev) return fail_msg(OMR_E_INVALID, "memory: gpu: device %d of %d", cfg->device, ndev);
  omr_store* s = new omr_store();
  s->cfg = *cfg;
  if (s->cfg.row_stride <= 0) s->cfg.row_stride = 1;
  s->dim = cfg->dim;
  s->pitch = ((cfg->dim + kBlockK - 1) / kBlockK) * kBlockK;
  s->cap = ((cfg->capacity + kTileRows - 1) / kTileRows) * kTileRows;
  if (s->cap > (int64_t)INT32_MAX - kTileRows) {
    delete s;
    return fail_msg(OMR_E_CAPACITY, "memory: a shard holds at most 2^31 rows");
  }
  // default: the deterministic error bound of the bf16 stage, valid for every dim <= 2048
  s->margin = cfg->candidate_margin > 0 ? cfg->candidate_margin : 4.5e-3f;
  if (const char* e = getenv("OMR_LEGACY_POST")) s->legacy_post = e[0] == '1';
  if (const char* e = getenv("OMR_PDL")) s->use_pdl = e[0] != '0';
  if (const char* e = getenv("OMR_POST_TIMING")) s->post_timing = e[0] == '1';
  if (const char* e = getenv("OMR_STREAM_LDG")) g_force_ldg_stream = e[0] == '1';
  int rc = s->open();
  if (rc) {
    s->destroy();
    delete s;
    return rc;
  }
  *out = s;
  return OMR_OK;
}

void omr_close(omr_store* s) {
  if (!s) return;
  s->destroy();
  delete s;
}

int64_t omr_count(const omr_store* s) { return s ? s->n : 0; }
int64_t omr_bitmap_words(const omr_store* s) { return s ? s->words() : 0; }

static void grow_host(omr_store* s, int64_t total) {
  const size_t t = (size_t)total;
  s->h_ws.resize(t); s->h_user.resize(t); s->h_agent.resize(t); s->h_entity.resize(t);
  s->h_kind.resize(t); s->h_purpose.resize(t); s->h_conf.resize(t); s->h_flags.resize(t);
  s->h_valid.resize(t); s->obs_id.resize(t);
}

static uint32_t entity_of(omr_store* s, const Key128& e) {
  auto it = s->entity_idx.find(e);
  if (it != s->entity_idx.end()) return it->second;
  const uint32_t idx = (uint32_t)s->entity_id.size();
  s->entity_id.push_back(e);
  s->entity_latest.push_back(-1);
  s->entity_rows.emplace_back();
  s->entity_idx.emplace(e, idx);
  return idx;
}

// A new embedded row becomes its entity's representative (the row UpdateEmbedding targets).
static void attach_row(omr_store* s, int64_t row, uint32_t ent) {
  s->h_entity[row] = ent;
  s->entity_rows[ent].push_back(row);
  const int64_t prev = s->entity_latest[ent];
  if (prev >= 0 && prev != row) s->h_flags[prev] &= (uint8_t)~kRowEntityRep;
  s->entity_latest[ent] = row;
  s->h_flags[row] |= kRowEntityRep;
}

int omr_bulk_load(omr_store* s, const omr_bulk* b) {
  if (!s || !b) return fail_msg(OMR_E_INVALID, "memory: omr_bulk_load: null argument");
  if (b->n <= 0) return OMR_OK;
  if (!b->vectors) return fail_msg(OMR_E_INVALID, "memory: omr_bulk_load: vectors is null");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  if (s->n + b->n > s->cfg.capacity) return fail_msg(OMR_E_CAPACITY, "memory: gpu store is full (%lld + %lld > %lld rows)",
                                                     (long long)s->n, (long long)b->n, (long long)s->cfg.capacity);
  auto table = [](Interner& in, const char* const* names, int cnt, const char* what, std::vector<uint32_t>* map) {
    map->resize((size_t)std::max(cnt, 0));
    for (int i = 0; i < cnt; i++) (*map)[i] = in.intern(names[i] ? names[i] : "");
    (void)what;
  };
  std::vector<uint32_t> mws, muser, magent, mkind, mpurpose;
  table(s->ws_names, b->workspaces, b->n_workspaces, "workspace", &mws);
  table(s->user_names, b->users, b->n_users, "user", &muser);
  table(s->agent_names, b->agents, b->n_agents, "agent", &magent);
  table(s->kind_names, b->kinds, b->n_kinds, "kind", &mkind);
  table(s->purpose_names, b->purposes, b->n_purposes, "purpose", &mpurpose);
  const uint32_t default_ws = mws.empty() ? s->ws_names.intern("default") : mws[0];
  const uint32_t default_kind = mkind.empty() ? s->kind_names.intern("") : mkind[0];
  const uint32_t default_purpose = mpurpose.empty() ? s->purpose_names.intern("") : mpurpose[0];
  const int64_t row0 = s->n;
  grow_host(s, row0 + b->n);
  for (int64_t i = 0; i < b->n; i++) {
    const int64_t r = row0 + i;
    auto pick = [&](const std::vector<uint32_t>& m, int64_t idx, uint32_t dflt, const char* what) -> int64_t {
      if (idx < 0 || idx >= (int64_t)m.size()) {
        fail_msg(OMR_E_INVALID, "memory: omr_bulk_load: row %lld %s index %lld outside table of %zu", (long long)i, what,
                 (long long)idx, m.size());
        return -1;
      }
      (void)dflt;
      return m[(size_t)idx];
    };
    int64_t v;
    if (b->workspace_idx) { if ((v = pick(mws, b->workspace_idx[i], 0, "workspace")) < 0) { grow_host(s, row0); return OMR_E_INVALID; } s->h_ws[r] = (uint32_t)v; }
    else s->h_ws[r] = default_ws;
    if (b->user_idx && b->user_idx[i] >= 0) { if ((v = pick(muser, b->user_idx[i], 0, "user")) < 0) { grow_host(s, row0); return OMR_E_INVALID; } s->h_user[r] = (uint32_t)v; }
    else s->h_user[r] = 0;
    if (b->agent_idx && b->agent_idx[i] >= 0) { if ((v = pick(magent, b->agent_idx[i], 0, "agent")) < 0) { grow_host(s, row0); return OMR_E_INVALID; } s->h_agent[r] = (uint32_t)v; }
    else s->h_agent[r] = 0;
    if (b->kind_idx) { if ((v = pick(mkind, b->kind_idx[i], 0, "kind")) < 0) { grow_host(s, row0); return OMR_E_INVALID; } s->h_kind[r] = (uint16_t)v; }
    else s->h_kind[r] = (uint16_t)default_kind;
    if (b->purpose_idx) { if ((v = pick(mpurpose, b->purpose_idx[i], 0, "purpose")) < 0) { grow_host(s, row0); return OMR_E_INVALID; } s->h_purpose[r] = (uint16_t)v; }
    else s->h_purpose[r] = (uint16_t)default_purpose;
    s->h_conf[r] = b->confidence ? b->confidence[i] : 1.0f;
    s->h_flags[r] = b->flags ? (uint8_t)(b->flags[i] & 0x0F) : (uint8_t)(OMR_ROW_HAS_EMBEDDING | OMR_ROW_MUTABLE);
    s->h_valid[r] = b->valid_until_us ? b->valid_until_us[i] : OMR_NO_EXPIRY;
    s->obs_id[r] = b->observation_ids ? key_of(b->observation_ids + 16 * i) : key_from_index(r);
  }
  // The vectors go in first, into rows beyond `n` that nothing can see yet: a NaN or an infinite element fails the
  // whole load, as the statement would fail in Postgres, and leaves no trace.
  int rc = s->append(b->n, b->vectors, row0);
  if (rc) {
    grow_host(s, row0);
    return rc;
  }
  // identity maps (after validation, so a rejected batch leaves no trace).  An observation id that is already
  // resident -- in the store or earlier in this batch -- is superseded by the new row: the old row is
  // hard-deleted (UpdateObservationEmbedding keeps one row per observation, store_write.go:580-600).
  for (int64_t i = 0; i < b->n; i++) {
    const int64_t r = row0 + i;
    auto old = s->obs_row.find(s->obs_id[r]);
    if (old != s->obs_row.end() && old->second != r) {
      const int64_t o = old->second;
      s->h_flags[o] = (uint8_t)((s->h_flags[o] | kRowDeleted) & ~OMR_ROW_HAS_EMBEDDING);
      s->release_row(o);
    }
    s->obs_row[s->obs_id[r]] = r;
    const Key128 e = b->entity_ids ? key_of(b->entity_ids + 16 * i) : s->obs_id[r];
    attach_row(s, r, entity_of(s, e));
  }
  s->n = row0 + b->n;
  rc = s->push_meta(0, s->n);  // earlier rows may have lost their EntityRep bit
  if (rc) return rc;
  s->live_dirty = true;
  return OMR_OK;
}

// Host half of an upsert: picks the row (in place for a resident observation id, else a reclaimed slot, else a new
// row at the end), fills the host mirrors and the identity maps.  `also` collects other rows whose flags changed.
static int upsert_host(omr_store* s, const omr_row* row, int64_t* row_out, std::vector<int64_t>* also) {
  if (!row->workspace_id || !row->workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  const Key128 ok = key_of(row->observation_id);
  int64_t r;
  auto it = s->obs_row.find(ok);
  const bool fresh = it == s->obs_row.end();
  bool appended = false;
  if (fresh) {
    r = s->recycle_row(also);  // the slot of a hard-deleted row, if one waits
    if (r < 0) {
      if (s->n + 1 > s->cfg.capacity)
        return fail_msg(OMR_E_CAPACITY, "memory: gpu store is full (%lld rows, none reclaimable)", (long long)s->n);
      r = s->n;
      grow_host(s, r + 1);
      appended = true;
    }
    s->obs_id[r] = ok;
    s->h_flags[r] = 0;
  } else {
    r = it->second;
  }
  s->h_ws[r] = s->ws_names.intern(row->workspace_id);
  s->h_user[r] = row->virtual_user_id ? s->user_names.intern(row->virtual_user_id) : 0;
  s->h_agent[r] = row->agent_id ? s->agent_names.intern(row->agent_id) : 0;
  s->h_kind[r] = (uint16_t)s->kind_names.intern(row->kind ? row->kind : "");
  s->h_purpose[r] = (uint16_t)s->purpose_names.intern(row->purpose ? row->purpose : "");
  s->h_conf[r] = row->confidence;
  s->h_flags[r] = (uint8_t)((s->h_flags[r] & kRowEntityRep) | (row->flags & 0x0F) | OMR_ROW_HAS_EMBEDDING);
  s->h_valid[r] = row->valid_until_us;
  if (fresh) {
    s->obs_row[ok] = r;
    const uint32_t ent = entity_of(s, key_of(row->entity_id));
    const int64_t prev = s->entity_latest[ent];
    attach_row(s, r, ent);
    if (appended) s->n = r + 1;
    if (prev >= 0 && prev != r) also->push_back(prev);
  }
  *row_out = r;
  return OMR_OK;
}

int omr_upsert(omr_store* s, const omr_row* row, const float* vec, int32_t dim) {
  if (!s || !row || !vec) return fail_msg(OMR_E_INVALID, "memory: omr_upsert: null argument");
  if (dim != s->dim) return fail_msg(OMR_E_DIM, "memory: embedding dimension mismatch: got %d, store has %d", dim, s->dim);
  if (!row->workspace_id || !row->workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  if (int frc = check_finite(vec, (size_t)dim)) return frc;
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  int64_t r;
  std::vector<int64_t> also;
  int rc = upsert_host(s, row, &r, &also);
  if (rc) return rc;
  if ((rc = s->push_meta(r, 1))) return rc;
  for (int64_t x : also)
    if ((rc = s->push_flag(x))) return rc;
  return s->append(1, vec, r);
}

int omr_upsert_batch(omr_store* s, const omr_row* rows, const float* vecs, int32_t n, int32_t dim) {
  if (!s || (n > 0 && (!rows || !vecs))) return fail_msg(OMR_E_INVALID, "memory: omr_upsert_batch: null argument");
  if (n <= 0) return OMR_OK;
  if (dim != s->dim) return fail_msg(OMR_E_DIM, "memory: embedding dimension mismatch: got %d, store has %d", dim, s->dim);
  for (int32_t i = 0; i < n; i++)
    if (!rows[i].workspace_id || !rows[i].workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  if (int frc = check_finite(vecs, (size_t)n * dim)) return frc;
  std::lock_guard<std::mutex> gate(s->join_gate);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  int rc;
  std::vector<int64_t> dst((size_t)n), also;
  for (int32_t i = 0; i < n; i++)
    if ((rc = upsert_host(s, &rows[i], &dst[(size_t)i], &also))) return rc;   // rows before i stay upserted, as n single calls would
  // an observation id given twice keeps its last vector: one writer per row in the ingest kernel
  std::unordered_map<int64_t, int32_t> last;
  for (int32_t i = 0; i < n; i++) last[dst[(size_t)i]] = i;
  std::vector<float> picked;
  const float* src = vecs;
  std::vector<int64_t> rows_u;
  if ((int32_t)last.size() != n) {
    picked.reserve(last.size() * (size_t)dim);
    for (int32_t i = 0; i < n; i++)
      if (last[dst[(size_t)i]] == i) {
        picked.insert(picked.end(), vecs + (size_t)i * dim, vecs + (size_t)(i + 1) * dim);
        rows_u.push_back(dst[(size_t)i]);
      }
    src = picked.data();
  } else {
    rows_u = dst;
  }
  // metadata: one copy per run of consecutive rows (new rows are consecutive), single flags for the rest
  std::vector<int64_t> sorted = rows_u;
  std::sort(sorted.begin(), sorted.end());
  for (size_t i = 0; i < sorted.size();) {
    size_t j = i + 1;
    while (j < sorted.size() && sorted[j] == sorted[j - 1] + 1) j++;
    if ((rc = s->push_meta(sorted[i], (int64_t)(j - i)))) return rc;
    i = j;
  }
  for (int64_t x : also)
    if (x < s->n && (rc = s->push_flag(x))) return rc;
  return s->append((int64_t)rows_u.size(), src, 0, rows_u.data());
}

int omr_compact(omr_store* s, int64_t* reclaimed) {
  if (!s) return fail_msg(OMR_E_INVALID, "memory: omr_compact: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  int64_t got = 0;
  const int rc = s->compact(&got);
  if (reclaimed) *reclaimed = got;
  return rc;
}

int64_t omr_free_slots(const omr_store* s) {
  if (!s) return 0;
  int64_t c = 0;
  for (int64_t r : s->free_rows)
    if (r < s->n && (s->h_flags[(size_t)r] & kRowFree)) c++;
  return c;
}

int omr_update_embedding(omr_store* s, const uint8_t entity_id[16], const float* vec, int32_t dim) {
  if (!s || !entity_id || !vec) return fail_msg(OMR_E_INVALID, "memory: omr_update_embedding: null argument");
  if (dim != s->dim) return fail_msg(OMR_E_DIM, "memory: embedding dimension mismatch: got %d, store has %d", dim, s->dim);
  if (int frc = check_finite(vec, (size_t)dim)) return frc;
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  auto it = s->entity_idx.find(key_of(entity_id));
  if (it == s->entity_idx.end() || s->entity_latest[it->second] < 0) return fail_msg(OMR_E_NOT_FOUND, "memory: not found");
  const int64_t r = s->entity_latest[it->second];
  s->h_flags[r] |= OMR_ROW_HAS_EMBEDDING;
  int rc = s->push_meta(r, 1);
  if (rc) return rc;
  return s->append(1, vec, r);
}

static int edit_rows(omr_store* s, const std::vector<int64_t>& rows, uint32_t set_mask, uint32_t clear_mask) {
  for (int64_t r : rows) {
    s->h_flags[r] = (uint8_t)((s->h_flags[r] | (set_mask & 0x4F)) & ~(clear_mask & 0x4F));
    if (set_mask & kRowDeleted) s->release_row(r);
    int rc = s->push_flag(r);
    if (rc) return rc;
  }
  return OMR_OK;
}

int omr_set_row_flags(omr_store* s, const uint8_t observation_id[16], uint32_t set_mask, uint32_t clear_mask) {
  if (!s || !observation_id) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  auto it = s->obs_row.find(key_of(observation_id));
  if (it == s->obs_row.end()) return fail_msg(OMR_E_NOT_FOUND, "memory: not found");
  return edit_rows(s, {it->second}, set_mask, clear_mask);
}

int omr_set_entity_flags(omr_store* s, const uint8_t entity_id[16], uint32_t set_mask, uint32_t clear_mask) {
  if (!s || !entity_id) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  auto it = s->entity_idx.find(key_of(entity_id));
  if (it == s->entity_idx.end()) return fail_msg(OMR_E_NOT_FOUND, "memory: not found");
  return edit_rows(s, s->entity_rows[it->second], set_mask, clear_mask);
}

int omr_delete_entity(omr_store* s, const uint8_t entity_id[16]) {
  return omr_set_entity_flags(s, entity_id, kRowDeleted, OMR_ROW_HAS_EMBEDDING);
}

int omr_delete_observation(omr_store* s, const uint8_t observation_id[16]) {
  return omr_set_row_flags(s, observation_id, kRowDeleted, OMR_ROW_HAS_EMBEDDING);
}

// Pushes the metadata of a set of rows: one range copy when the set is large, row by row otherwise.
static int push_rows(omr_store* s, const std::vector<int64_t>& rows) {
  if (rows.empty()) return OMR_OK;
  if (rows.size() > 64) {
    const auto mm = std::minmax_element(rows.begin(), rows.end());
    return s->push_meta(*mm.first, *mm.second - *mm.first + 1);
  }
  for (int64_t r : rows)
    if (int rc = s->push_meta(r, 1)) return rc;
  return OMR_OK;
}

static bool entity_is_live(const omr_store* s, uint32_t ent) {
  for (int64_t r : s->entity_rows[ent])
    if (!(s->h_flags[r] & kRowDeleted)) return true;
  return false;
}

int omr_delete_scope(omr_store* s, const char* workspace_id, const char* virtual_user_id, int64_t limit,
                     int64_t* deleted) {
  if (deleted) *deleted = 0;
  if (!s) return fail_msg(OMR_E_INVALID, "memory: null argument");
  if (!workspace_id || !workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  uint32_t ws = 0, user = 0;
  if (!s->ws_names.lookup(workspace_id, &ws)) return OMR_OK;  // nothing stored under this workspace
  const bool by_user = virtual_user_id && virtual_user_id[0];
  if (by_user && !s->user_names.lookup(virtual_user_id, &user)) return OMR_OK;
  std::vector<int64_t> touched;
  int64_t count = 0;
  for (uint32_t e = 0; e < (uint32_t)s->entity_rows.size(); e++) {
    if (limit > 0 && count >= limit) break;
    const auto& rows = s->entity_rows[e];
    if (rows.empty() || !entity_is_live(s, e)) continue;
    const int64_t r0 = rows[0];  // entity-level columns are the same on every row of the entity
    if (s->h_ws[r0] != ws || (by_user && s->h_user[r0] != user)) continue;
    for (int64_t r : rows) {
      s->h_flags[r] = (uint8_t)((s->h_flags[r] | kRowDeleted) & ~OMR_ROW_HAS_EMBEDDING);
      s->release_row(r);
      touched.push_back(r);
    }
    count++;
  }
  if (deleted) *deleted = count;
  if (touched.size() > 64) {
    const auto mm = std::minmax_element(touched.begin(), touched.end());
    OMR_CUDA_OK(cudaMemcpy(s->d_flags + *mm.first, s->h_flags.data() + *mm.first, (size_t)(*mm.second - *mm.first + 1),
                           cudaMemcpyHostToDevice));
    s->live_dirty = true;
    return OMR_OK;
  }
  for (int64_t r : touched)
    if (int rc = s->push_flag(r)) return rc;
  return OMR_OK;
}

int omr_rescope(omr_store* s, const uint8_t* observation_ids, int32_t n, const char* virtual_user_id,
                const char* agent_id) {
  if (!s || (n > 0 && !observation_ids)) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  const uint32_t user = (virtual_user_id && virtual_user_id[0]) ? s->user_names.intern(virtual_user_id) : 0;
  const uint32_t agent = (agent_id && agent_id[0]) ? s->agent_names.intern(agent_id) : 0;
  std::vector<int64_t> touched;
  std::vector<uint32_t> done;
  for (int32_t i = 0; i < n; i++) {
    auto it = s->obs_row.find(key_of(observation_ids + 16 * (size_t)i));
    if (it == s->obs_row.end() || (s->h_flags[it->second] & kRowDeleted)) continue;
    const uint32_t ent = s->h_entity[it->second];
    if (std::find(done.begin(), done.end(), ent) != done.end()) continue;
    done.push_back(ent);
    for (int64_t r : s->entity_rows[ent]) {
      s->h_user[r] = user;
      s->h_agent[r] = agent;
      touched.push_back(r);
    }
  }
  return push_rows(s, touched);
}

int omr_merge_entities(omr_store* s, const uint8_t canonical_id[16], const uint8_t* merge_ids, int32_t n) {
  if (!s || !canonical_id || (n > 0 && !merge_ids)) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  const uint32_t canon = entity_of(s, key_of(canonical_id));
  // entity-level columns of the canonical entity, when it has a resident row to read them from
  int64_t src = -1;
  for (int64_t r : s->entity_rows[canon])
    if (!(s->h_flags[r] & kRowDeleted)) { src = r; break; }
  const uint8_t entity_bits = (uint8_t)(OMR_ROW_FORGOTTEN | OMR_ROW_MUTABLE);
  std::vector<int64_t> touched;
  for (int32_t i = 0; i < n; i++) {
    auto it = s->entity_idx.find(key_of(merge_ids + 16 * (size_t)i));
    if (it == s->entity_idx.end() || it->second == canon) continue;
    const uint32_t ent = it->second;
    for (int64_t r : s->entity_rows[ent]) {
      s->h_entity[r] = canon;
      s->entity_rows[canon].push_back(r);
      if (src >= 0) {
        s->h_ws[r] = s->h_ws[src]; s->h_user[r] = s->h_user[src]; s->h_agent[r] = s->h_agent[src];
        s->h_kind[r] = s->h_kind[src]; s->h_purpose[r] = s->h_purpose[src];
        s->h_flags[r] = (uint8_t)((s->h_flags[r] & ~entity_bits) | (s->h_flags[src] & entity_bits));
      }
      touched.push_back(r);
    }
    s->entity_rows[ent].clear();
    s->entity_latest[ent] = -1;
  }
  if (touched.empty()) return OMR_OK;
  // the representative (the row UpdateEmbedding targets, the row the self-join reads) is the newest
  // live embedded row of the merged entity
  int64_t rep = -1;
  for (int64_t r : s->entity_rows[canon]) {
    if (s->h_flags[r] & kRowEntityRep) touched.push_back(r);
    s->h_flags[r] &= (uint8_t)~kRowEntityRep;
    if (!(s->h_flags[r] & kRowDeleted) && r > rep) rep = r;
  }
  if (rep < 0 && !s->entity_rows[canon].empty())
    rep = *std::max_element(s->entity_rows[canon].begin(), s->entity_rows[canon].end());
  if (rep >= 0) {
    s->h_flags[rep] |= kRowEntityRep;
    touched.push_back(rep);
  }
  s->entity_latest[canon] = rep;
  return push_rows(s, touched);
}

int omr_clear(omr_store* s) {
  if (!s) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  s->n = 0;
  grow_host(s, 0);
  s->entity_id.clear(); s->entity_latest.clear(); s->entity_rows.clear();
  s->obs_row.clear(); s->entity_idx.clear();
  s->free_rows.clear();
  s->live_dirty = true;
  return OMR_OK;
}

int omr_set_valid_until(omr_store* s, const uint8_t observation_id[16], int64_t valid_until_us) {
  if (!s || !observation_id) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  auto it = s->obs_row.find(key_of(observation_id));
  if (it == s->obs_row.end()) return fail_msg(OMR_E_NOT_FOUND, "memory: not found");
  s->h_valid[it->second] = valid_until_us;
  OMR_CUDA_OK(cudaMemcpy(s->d_valid + it->second, &valid_until_us, 8, cudaMemcpyHostToDevice));
  return OMR_OK;
}

int omr_set_confidence(omr_store* s, const uint8_t observation_id[16], float confidence) {
  if (!s || !observation_id) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> gate(s->join_gate);   // rows do not change under a running self-join
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  auto it = s->obs_row.find(key_of(observation_id));
  if (it == s->obs_row.end()) return fail_msg(OMR_E_NOT_FOUND, "memory: not found");
  s->h_conf[it->second] = confidence;
  OMR_CUDA_OK(cudaMemcpy(s->d_conf + it->second, &confidence, 4, cudaMemcpyHostToDevice));
  return OMR_OK;
}

int omr_search(omr_store* s, const float* queries, int32_t nq, int32_t dim, const omr_filter* filters,
               int32_t n_filters, const int32_t* query_filter, int32_t k, omr_hit* hits, int32_t* counts) {
  if (!s || !queries || !hits || !counts) return fail_msg(OMR_E_INVALID, "memory: omr_search: null argument");
  if (dim != s->dim)
    return fail_msg(OMR_E_DIM, "memory: similar observations: different vector dimensions %d and %d", s->dim, dim);
  if (k <= 0) k = 5;  // DefaultDuplicateCandidateLimit, internal/memory/types.go:137-141
  if (nq <= 0) return OMR_OK;
  omr_store::HostReq r{queries, nq, filters, n_filters, query_filter, k, hits, counts, nullptr, nullptr};
  return s->submit(r);
}

int omr_search_rows(omr_store* s, const float* queries, int32_t nq, int32_t dim, const omr_filter* filters,
                    int32_t n_filters, const int32_t* query_filter, int32_t k, int64_t* rows, double* distances,
                    int32_t* counts) {
  if (!s || !queries || !rows || !distances || !counts) return fail_msg(OMR_E_INVALID, "memory: omr_search_rows: null argument");
  if (dim != s->dim)
    return fail_msg(OMR_E_DIM, "memory: similar observations: different vector dimensions %d and %d", s->dim, dim);
  if (k <= 0) k = 5;
  if (nq <= 0) return OMR_OK;
  omr_store::HostReq r{queries, nq, filters, n_filters, query_filter, k, nullptr, counts, rows, distances};
  return s->submit(r);
}

// ---- host buffers, pipelined ------------------------------------------------------------------

int omr_search_rows_begin(omr_store* s, const float* queries, int32_t nq, int32_t dim, const omr_filter* filters,
                          int32_t n_filters, const int32_t* query_filter, int32_t k, int32_t* ticket) {
  if (!s || !queries || !ticket) return fail_msg(OMR_E_INVALID, "memory: omr_search_rows_begin: null argument");
  if (dim != s->dim)
    return fail_msg(OMR_E_DIM, "memory: similar observations: different vector dimensions %d and %d", s->dim, dim);
  if (nq <= 0) return fail_msg(OMR_E_INVALID, "memory: omr_search_rows_begin: empty batch");
  if (n_filters <= 0 || !filters) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  if (k <= 0) k = 5;
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  int idx = -1;
  for (int i = 1; i < omr_store::kLanes; i++)
    if (!s->lanes[i].busy) { idx = i; break; }
  if (idx < 0) return fail_msg(OMR_E_CAPACITY, "memory: %d batches are already in flight; finish one first", omr_store::kLanes - 1);
  omr_store::Lane& L = s->lanes[idx];
  int rc;
  const size_t qbytes = (size_t)nq * s->dim * 4;
  const size_t fbytes = (size_t)n_filters * sizeof(DeviceFilter);
  // pinned staging: filters | max_dist | query_group | (queries, unless the caller's buffer is page-locked)
  const size_t off_m = (fbytes + 15) & ~size_t(15), off_g = off_m + (size_t)nq * 8, off_q = (off_g + (size_t)nq * 4 + 15) & ~size_t(15);
  if ((rc = L.hin.reserve(off_q + qbytes))) return rc;
  DeviceFilter* df = reinterpret_cast<DeviceFilter*>(L.hin.p);
  double* pm = reinterpret_cast<double*>(L.hin.p + off_m);
  int32_t* pg = reinterpret_cast<int32_t*>(L.hin.p + off_g);
  std::vector<double> fmax((size_t)n_filters);
  for (int i = 0; i < n_filters; i++)
    if ((rc = s->to_device_filter(filters[i], &df[i], &fmax[i]))) return rc;
  for (int i = 0; i < nq; i++) {
    const int g = query_filter ? query_filter[i] : 0;
    if (g < 0 || g >= n_filters) return fail_msg(OMR_E_INVALID, "memory: query %d names filter %d of %d", i, g, n_filters);
    pg[i] = g;
    pm[i] = fmax[g];
  }
  const float* qsrc = reinterpret_cast<const float*>(L.hin.p + off_q);
  {
    cudaPointerAttributes attr{};
    if (cudaPointerGetAttributes(&attr, queries) == cudaSuccess && attr.type == cudaMemoryTypeHost) qsrc = queries;
    else { cudaGetLastError(); memcpy(L.hin.p + off_q, queries, qbytes); }
  }
  if ((rc = L.hq.reserve((size_t)nq * s->dim))) return rc;
  if ((rc = L.hgroup.reserve((size_t)nq))) return rc;
  if ((rc = L.hmaxd.reserve((size_t)nq))) return rc;
  if ((rc = L.hfilters.reserve((size_t)n_filters))) return rc;
  if ((rc = L.hbitmaps.reserve((size_t)n_filters * s->words()))) return rc;
  if ((rc = L.hrows.reserve((size_t)nq * k))) return rc;
  if ((rc = L.hdist.reserve((size_t)nq * k))) return rc;
  if ((rc = L.hcount.reserve((size_t)nq))) return rc;
  if ((rc = L.hout.reserve((size_t)nq * k * 16 + (size_t)nq * 4))) return rc;
  cudaStream_t st = L.stream;
  OMR_CUDA_OK(cudaMemcpyAsync(L.hq.p, qsrc, qbytes, cudaMemcpyHostToDevice, st));
  OMR_CUDA_OK(cudaMemcpyAsync(L.hgroup.p, pg, (size_t)nq * 4, cudaMemcpyHostToDevice, st));
  OMR_CUDA_OK(cudaMemcpyAsync(L.hmaxd.p, pm, (size_t)nq * 8, cudaMemcpyHostToDevice, st));
  OMR_CUDA_OK(cudaMemcpyAsync(L.hfilters.p, df, fbytes, cudaMemcpyHostToDevice, st));
  if ((rc = launch_build_bitmaps(s->meta(), s->n, (int)s->words(), L.hfilters.p, n_filters, L.hbitmaps.p, st))) return rc;
  s->st_launches++;
  bool fused = true;
  if ((rc = s->enqueue(L, L.hq.p, nq, L.hbitmaps.p, L.hgroup.p, L.hmaxd.p, k, L.hrows.p, L.hdist.p, L.hcount.p, st, &fused)))
    return rc;
  L.immediate = false;
  if (!fused || L.nfail_known >= 0) {
    if ((rc = s->complete(L, st))) return rc;
    L.immediate = true;
  }
  int64_t* orow = reinterpret_cast<int64_t*>(L.hout.p);
  OMR_CUDA_OK(cudaMemcpyAsync(orow, L.hrows.p, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaMemcpyAsync(L.hout.p + (size_t)nq * k * 8, L.hdist.p, (size_t)nq * k * 8, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaMemcpyAsync(L.hout.p + (size_t)nq * k * 16, L.hcount.p, (size_t)nq * 4, cudaMemcpyDeviceToHost, st));
  L.hnq = nq;
  L.hk = k;
  L.host_batch = true;
  L.busy = true;
  *ticket = idx;
  return OMR_OK;
}

int omr_search_rows_finish(omr_store* s, int32_t ticket, int64_t* rows, double* distances, int32_t* counts) {
  if (!s || !rows || !distances || !counts || ticket < 1 || ticket >= omr_store::kLanes)
    return fail_msg(OMR_E_INVALID, "memory: omr_search_rows_finish: bad argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  omr_store::Lane& L = s->lanes[ticket];
  if (!L.busy || !L.host_batch) return fail_msg(OMR_E_INVALID, "memory: ticket %d is not a host batch in flight", ticket);
  int rc = OMR_OK;
  const size_t nk = (size_t)L.hnq * L.hk;
  if (!L.immediate) {
    rc = s->complete(L, L.stream);
    if (rc == OMR_OK && L.nfail_seen > 0) {   // the exact fallback rewrote some answers after the first read-back
      cudaMemcpyAsync(L.hout.p, L.hrows.p, nk * 8, cudaMemcpyDeviceToHost, L.stream);
      cudaMemcpyAsync(L.hout.p + nk * 8, L.hdist.p, nk * 8, cudaMemcpyDeviceToHost, L.stream);
      cudaMemcpyAsync(L.hout.p + nk * 16, L.hcount.p, (size_t)L.hnq * 4, cudaMemcpyDeviceToHost, L.stream);
    }
  }
  if (rc == OMR_OK) {
    cudaError_t e = cudaStreamSynchronize(L.stream);
    if (e != cudaSuccess) rc = fail_cuda(e, "cudaStreamSynchronize", __FILE__, __LINE__);
  }
  if (rc == OMR_OK) {
    memcpy(rows, L.hout.p, nk * 8);
    memcpy(distances, L.hout.p + nk * 8, nk * 8);
    memcpy(counts, L.hout.p + nk * 16, (size_t)L.hnq * 4);
  }
  L.busy = false;
  L.host_batch = false;
  return rc;
}

int omr_cosine_ranked(omr_store* s, const float* query, int32_t dim, const omr_filter* filter, int32_t fanout,
                      omr_hit* hits, int32_t* count) {
  if (!s || !query || !filter || !hits || !count) return fail_msg(OMR_E_INVALID, "memory: omr_cosine_ranked: null argument");
  if (dim != s->dim)
    return fail_msg(OMR_E_DIM, "memory: hybrid retrieve: different vector dimensions %d and %d", s->dim, dim);
  if (fanout <= 0) return fail_msg(OMR_E_INVALID, "memory: fanout must be positive");
  const int over = 4 * fanout;  // `LIMIT $6 * 4`, store.go:260
  if (over > 4096) return fail_msg(OMR_E_INVALID, "memory: fanout %d exceeds 1024", fanout);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  std::vector<omr_hit> ann((size_t)over);
  int32_t got = 0;
  omr_filter f = *filter;
  f.max_distance = NAN;  // the recall arm has no distance cutoff
  int rc = s->host_search(query, 1, &f, 1, nullptr, over, ann.data(), &got);
  if (rc) return rc;
  // cosine: DISTINCT ON (entity_id) nearest observation; cosine_ranked: ORDER BY cos_dist LIMIT fanout.
  std::unordered_map<Key128, bool, Key128Hash> seen;
  int out = 0;
  for (int i = 0; i < got && out < fanout; i++) {
    const Key128 e = key_of(ann[i].entity_id);
    if (seen.emplace(e, true).second) hits[out++] = ann[i];
  }
  *count = out;
  return OMR_OK;
}

int omr_device_build_bitmaps(omr_store* s, const omr_filter* filters, int32_t n_filters, uint32_t* d_bitmaps,
                             void* stream) {
  if (!s || !filters || !d_bitmaps || n_filters <= 0) return fail_msg(OMR_E_INVALID, "memory: omr_device_build_bitmaps: bad argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  std::vector<DeviceFilter> df((size_t)n_filters);
  double unused;
  int rc;
  for (int i = 0; i < n_filters; i++)
    if ((rc = s->to_device_filter(filters[i], &df[i], &unused))) return rc;
  if ((rc = s->w_filters.reserve((size_t)std::max(64, n_filters + 1)))) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  OMR_CUDA_OK(cudaMemcpyAsync(s->w_filters.p, df.data(), (size_t)n_filters * sizeof(DeviceFilter), cudaMemcpyHostToDevice, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  rc = launch_build_bitmaps(s->meta(), s->n, (int)s->words(), s->w_filters.p, n_filters, d_bitmaps, st);
  s->st_launches++;
  return rc;
}

int omr_device_search(omr_store* s, const float* d_queries, int32_t nq, const uint32_t* d_bitmaps, int32_t n_groups,
                      const int32_t* d_query_group, const double* d_max_distance, int32_t k, int64_t* d_rows,
                      double* d_dist, int32_t* d_count, void* stream) {
  if (!s || !d_queries || !d_rows || !d_dist || !d_count) return fail_msg(OMR_E_INVALID, "memory: omr_device_search: null argument");
  (void)n_groups;
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  return s->device_search(d_queries, nq, d_bitmaps, d_query_group, d_max_distance, k, d_rows, d_dist, d_count,
                          (cudaStream_t)stream);
}

int omr_device_search_begin(omr_store* s, const float* d_queries, int32_t nq, const uint32_t* d_bitmaps,
                            int32_t n_groups, const int32_t* d_query_group, const double* d_max_distance, int32_t k,
                            int64_t* d_rows, double* d_dist, int32_t* d_count, void* stream, int32_t* ticket) {
  if (!s || !d_queries || !d_rows || !d_dist || !d_count || !ticket)
    return fail_msg(OMR_E_INVALID, "memory: omr_device_search_begin: null argument");
  (void)n_groups;
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  int idx = -1;
  for (int i = 1; i < omr_store::kLanes; i++)
    if (!s->lanes[i].busy) { idx = i; break; }
  if (idx < 0) return fail_msg(OMR_E_CAPACITY, "memory: %d batches are already in flight; finish one first", omr_store::kLanes - 1);
  omr_store::Lane& L = s->lanes[idx];
  cudaStream_t caller = (cudaStream_t)stream;
  // the lane's stream picks up after whatever the caller already enqueued (its inputs)
  OMR_CUDA_OK(cudaEventRecord(L.ev_in, caller));
  OMR_CUDA_OK(cudaStreamWaitEvent(L.stream, L.ev_in, 0));
  bool fused = true;
  int rc = s->enqueue(L, d_queries, nq, d_bitmaps, d_query_group, d_max_distance, k, d_rows, d_dist, d_count, L.stream,
                      &fused);
  if (rc) return rc;
  L.caller = caller;
  L.immediate = false;
  if (!fused) {
    // the unfused chain and the exact-only path share scratch between lanes: run them to completion now
    if ((rc = s->complete(L, L.stream))) return rc;
    OMR_CUDA_OK(cudaStreamSynchronize(L.stream));
    L.immediate = true;
  } else {
    OMR_CUDA_OK(cudaEventRecord(L.ev_done, L.stream));
  }
  L.busy = true;
  *ticket = idx;
  return OMR_OK;
}

int omr_device_search_finish(omr_store* s, int32_t ticket) {
  if (!s || ticket < 1 || ticket >= omr_store::kLanes) return fail_msg(OMR_E_INVALID, "memory: omr_device_search_finish: bad ticket %d", ticket);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  omr_store::Lane& L = s->lanes[ticket];
  if (!L.busy || L.host_batch || L.xslot >= 0) return fail_msg(OMR_E_INVALID, "memory: ticket %d is not in flight", ticket);
  int rc = OMR_OK;
  if (!L.immediate) {
    rc = s->complete(L, L.stream);   // waits for the lane, runs the exact fallback if a certificate failed
    if (rc == OMR_OK) {
      // later work on the caller's stream is ordered after the lane's results
      cudaError_t e = cudaEventRecord(L.ev_done, L.stream);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(L.caller, L.ev_done, 0);
      if (e != cudaSuccess) rc = fail_cuda(e, "event hand-off", __FILE__, __LINE__);
    }
  }
  L.busy = false;
  return rc;
}

// ---- fused exchange over peer memory ---------------------------------------------------------

int omr_exchange_alloc(omr_store* s, int32_t world, int32_t nq_max, int32_t k_max, int32_t slots, void** d_buf,
                       int64_t* bytes) {
  if (!s || !d_buf || !bytes) return fail_msg(OMR_E_INVALID, "memory: omr_exchange_alloc: null argument");
  if (world < 1 || world > kMaxWorld || nq_max < 1 || k_max < 1 || slots < 1 || slots > 64)
    return fail_msg(OMR_E_INVALID, "memory: omr_exchange_alloc: world 1..%d, slots 1..64", kMaxWorld);
  // a slot must not be rewritten before every rank has merged it: 2 x (batches in flight) + 1
  if (world > 1 && slots < 2 * (omr_store::kLanes - 1) + 1)
    return fail_msg(OMR_E_INVALID, "memory: omr_exchange_alloc: at least %d slots (2 x batches in flight + 1), got %d",
                    2 * (omr_store::kLanes - 1) + 1, slots);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (s->xbuf) return fail_msg(OMR_E_INVALID, "memory: the exchange buffer of this store is already allocated");
  ExchangeView v;
  v.world = world; v.slots = slots; v.rank = 0;
  v.slot_stride = (int64_t)2 * nq_max * k_max * 8;
  const size_t total = v.total_bytes();
  OMR_CUDA_OK(cudaMalloc(&s->xbuf, total));
  OMR_CUDA_OK(cudaMemset(s->xbuf, 0, total));   // flags start at epoch 0; the first batch of a slot is epoch 1
  s->xv = v;
  s->xnq_max = nq_max; s->xk_max = k_max;
  s->xslot_epoch.assign((size_t)slots, 0ull);
  int rc;
  if ((rc = s->xwords.reserve(16 + 16 * omr_store::kLanes))) return rc;   // [0] error word, then 16 block counters per lane
  OMR_CUDA_OK(cudaMemset(s->xwords.p, 0, (16 + 16 * omr_store::kLanes) * sizeof(unsigned int)));
  if ((rc = s->xerr_pin.reserve(4))) return rc;
  s->xerr_pin.p[0] = 0;
  *d_buf = s->xbuf;
  *bytes = (int64_t)total;
  return OMR_OK;
}

int omr_ipc_export(const void* d_ptr, uint8_t handle[64]) {
  if (!d_ptr || !handle) return fail_msg(OMR_E_INVALID, "memory: omr_ipc_export: null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  cudaIpcMemHandle_t h;
  OMR_CUDA_OK(cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr)));
  memcpy(handle, &h, 64);
  return OMR_OK;
}

int omr_exchange_setup(omr_store* s, int32_t rank, int32_t world, const uint8_t* handles /*[world][64]*/) {
  if (!s || !handles) return fail_msg(OMR_E_INVALID, "memory: omr_exchange_setup: null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (!s->xbuf || world != s->xv.world || rank < 0 || rank >= world)
    return fail_msg(OMR_E_INVALID, "memory: omr_exchange_setup: allocate first, with the same world size");
  s->xv.rank = rank;
  for (int p = 0; p < world; p++) {
    if (p == rank) {
      s->xv.peers[p] = s->xbuf;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, handles + 64 * p, 64);
    void* ptr = nullptr;
    OMR_CUDA_OK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    s->xv.peers[p] = ptr;
    s->xopened.push_back(ptr);
  }
  return OMR_OK;
}

int omr_device_search_exchange_begin(omr_store* s, const float* d_queries, int32_t nq, const uint32_t* d_bitmaps,
                                     int32_t n_groups, const int32_t* d_query_group, const double* d_max_distance,
                                     int32_t k, int32_t slot, void* stream, int32_t* ticket) {
  if (!s || !d_queries || !ticket) return fail_msg(OMR_E_INVALID, "memory: omr_device_search_exchange_begin: null argument");
  (void)n_groups;
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (!s->xbuf || s->xv.peers[s->xv.world - 1] == nullptr) return fail_msg(OMR_E_INVALID, "memory: call omr_exchange_setup first");
  if (nq < 1 || nq > s->xnq_max || k < 1 || k > s->xk_max || slot < 0 || slot >= s->xv.slots)
    return fail_msg(OMR_E_CAPACITY, "memory: batch (%d queries, k=%d, slot %d) exceeds the exchange buffer (%d, %d, %d slots)",
                    nq, k, slot, s->xnq_max, s->xk_max, s->xv.slots);
  if (s->xerr_pin.p[0]) return fail_msg(OMR_E_CUDA, "memory: gpu: a peer never published its answer (exchange timed out)");
  int idx = -1;
  for (int i = 1; i < omr_store::kLanes; i++)
    if (!s->lanes[i].busy) { idx = i; break; }
  if (idx < 0) return fail_msg(OMR_E_CAPACITY, "memory: %d batches are already in flight; finish one first", omr_store::kLanes - 1);
  omr_store::Lane& L = s->lanes[idx];
  int rc;
  if ((rc = L.xlocal.reserve((size_t)2 * nq * k))) return rc;
  if ((rc = L.xcount.reserve((size_t)nq))) return rc;
  cudaStream_t caller = (cudaStream_t)stream;
  OMR_CUDA_OK(cudaEventRecord(L.ev_in, caller));
  OMR_CUDA_OK(cudaStreamWaitEvent(L.stream, L.ev_in, 0));
  bool fused = true;
  int64_t* lrows = L.xlocal.p;
  double* ldist = reinterpret_cast<double*>(L.xlocal.p + (size_t)nq * k);
  rc = s->enqueue(L, d_queries, nq, d_bitmaps, d_query_group, d_max_distance, k, lrows, ldist, L.xcount.p, L.stream, &fused);
  if (rc) return rc;
  L.caller = caller;
  L.xslot = slot;
  L.xnq = nq;
  L.xk = k;
  L.xepoch = ++s->xslot_epoch[(size_t)slot];
  L.immediate = false;
  if (!fused || L.nfail_known >= 0) {
    // unfused chain / exact-only path / empty store: finish the local answer now; _finish publishes it
    if ((rc = s->complete(L, L.stream))) return rc;
    L.immediate = true;
  } else {
    // publish straight from the device unless a certificate failed (then _finish publishes after the fallback)
    if ((rc = launch_exchange_scatter(s->xv, L.xlocal.p, nq, k, slot, L.xepoch, L.counters.p + 1,
                                      s->xwords.p + 16 + 16 * idx, L.stream)))
      return rc;
    s->st_launches++;
  }
  L.busy = true;
  *ticket = idx;
  return OMR_OK;
}

int omr_device_search_exchange_finish(omr_store* s, int32_t ticket, int64_t* d_rows, double* d_dist, int32_t* d_count) {
  if (!s || !d_rows || !d_dist || !d_count || ticket < 1 || ticket >= omr_store::kLanes)
    return fail_msg(OMR_E_INVALID, "memory: omr_device_search_exchange_finish: bad argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  omr_store::Lane& L = s->lanes[ticket];
  if (!L.busy || L.xslot < 0) return fail_msg(OMR_E_INVALID, "memory: ticket %d is not an exchange batch in flight", ticket);
  int rc = OMR_OK;
  bool publish = L.immediate;
  if (!L.immediate) {
    rc = s->complete(L, L.stream);
    publish = L.nfail_seen > 0;   // the device held its flag back: publish the repaired answer
  }
  if (rc == OMR_OK) {
    if (publish) {
      rc = launch_exchange_scatter(s->xv, L.xlocal.p, L.xnq, L.xk, L.xslot, L.xepoch, nullptr,
                                   s->xwords.p + 16 + 16 * ticket, L.stream);
      s->st_launches++;
    }
    if (rc == OMR_OK) {
      rc = launch_exchange_merge(s->xv, L.xslot, L.xepoch, L.xnq, L.xk, d_rows, d_dist, d_count, s->xwords.p, L.stream);
      s->st_launches++;
    }
    if (rc == OMR_OK) {
      cudaError_t e = cudaMemcpyAsync(s->xerr_pin.p, s->xwords.p, sizeof(unsigned int), cudaMemcpyDeviceToHost, L.stream);
      if (e == cudaSuccess) e = cudaEventRecord(L.ev_done, L.stream);
      if (e == cudaSuccess) e = cudaStreamWaitEvent(L.caller, L.ev_done, 0);
      if (e != cudaSuccess) rc = fail_cuda(e, "exchange hand-off", __FILE__, __LINE__);
    }
  }
  L.busy = false;
  L.xslot = -1;
  return rc;
}

int omr_device_merge_topk(omr_store* s, const int64_t* d_in_rows, const double* d_in_dist, int32_t n_shards,
                          int32_t nq, int32_t k, int64_t shard_stride, int64_t* d_rows, double* d_dist, int32_t* d_count, void* stream) {
  if (!s || !d_in_rows || !d_in_dist || !d_rows || !d_dist || !d_count) return fail_msg(OMR_E_INVALID, "memory: omr_device_merge_topk: null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  s->st_launches++;
  return launch_merge_topk(d_in_rows, d_in_dist, n_shards, nq, k, shard_stride, d_rows, d_dist, d_count,
                           (cudaStream_t)stream);
}

// The join proper: member scan, tensor-core candidate pass (chunked launches, cancellable between chunks),
// exact rescore + row-pair rules + `ORDER BY similarity DESC LIMIT n` on the device.  Returns the selected pairs
// as (canonical row, dupe row) of THIS shard with their exact similarities, best first.  Caller holds join_gate.
int omr_store::selfjoin(const omr_selfjoin_opts* o, std::vector<PairCand>* out_pairs, std::vector<double>* out_sims) {
  out_pairs->clear();
  out_sims->clear();
  uint32_t ws = 0;
  int64_t n_rows = 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    cudaSetDevice(cfg.device);
    if (int drc = drain()) return drc;   // only for what earlier calls left on the shared stream
    if (!ws_names.lookup(o->workspace_id, &ws) || n == 0) return OMR_OK;
    n_rows = n;
    OMR_CUDA_OK(cudaStreamSynchronize(stream));   // pending metadata pushes are visible to jstream
  }
  cudaSetDevice(cfg.device);
  cudaStream_t st = jstream;
  int rc;
  const int part_count = o->part_count > 0 ? o->part_count : 1;
  const int part_index = o->part_index;
  if (part_index < 0 || part_index >= part_count) return fail_msg(OMR_E_INVALID, "memory: self-join part %d of %d", part_index, part_count);
  auto cancelled = [&]() { return o->cancel && *o->cancel != 0; };

  // ---- members: one row per entity of the workspace that carries a usable embedding (device scan)
  const int blocks = join_scan_blocks(n_rows);
  if ((rc = w_member.reserve((size_t)((n_rows + 31) / 32) + 8))) return rc;
  if ((rc = w_jcounts.reserve((size_t)blocks))) return rc;
  if ((rc = w_joffsets.reserve((size_t)blocks))) return rc;
  if ((rc = w_jtotal.reserve(1))) return rc;
  if ((rc = jpin.reserve(8))) return rc;
  if ((rc = launch_join_member_scan(meta(), n_rows, ws, o->require_mutable, w_member.p, w_jcounts.p, w_joffsets.p,
                                    w_jtotal.p, st)))
    return rc;
  st_launches += 2;
  OMR_CUDA_OK(cudaMemcpyAsync(jpin.p, w_jtotal.p, 8, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  const int64_t m = (int64_t)jpin.p[0];
  if (m < 2) return OMR_OK;
  const bool identity = m == n_rows;
  const int64_t m_pad = ((m + kTileRows - 1) / kTileRows) * kTileRows;
  if ((rc = w_idx.reserve((size_t)m))) return rc;
  if ((rc = w_kind_c.reserve((size_t)m_pad))) return rc;
  if ((rc = launch_join_member_scatter(w_member.p, n_rows, w_joffsets.p, d_kind, w_idx.p, w_kind_c.p, st))) return rc;
  st_launches++;
  CUtensorMap ta, tb;
  if (identity) {
    ta = tmap_rows_a;
    tb = tmap_rows_b;
  } else {
    if ((rc = w_gather.reserve((size_t)m_pad * pitch))) return rc;
    OMR_CUDA_OK(cudaMemsetAsync(w_gather.p + (size_t)m * pitch, 0, (size_t)(m_pad - m) * pitch * 2, st));
    if ((rc = launch_gather_rows(view(), w_idx.p, m, w_gather.p, st))) return rc;
    st_launches++;
    if ((rc = make_tmap(&ta, w_gather.p, (uint64_t)m_pad, kQueryTile))) return rc;
    if ((rc = make_tmap(&tb, w_gather.p, (uint64_t)m_pad, kTileRows))) return rc;
  }
  if ((rc = w_jbits.reserve((size_t)(m_pad / 32)))) return rc;
  if ((rc = launch_compact_member_bits(m, m_pad / 32, w_jbits.p, st))) return rc;
  st_launches++;

  // ---- candidate pass
  const double floor_d = (double)o->similarity_floor;  // float32 widened once, as Postgres does
  // Extra candidates cost little here, so the join always uses at least the deterministic bound of the bf16
  // stage (2^-8 from the two roundings of unit vectors, plus fp32 accumulation over <= 2048 terms): the pair set
  // is exact under any rounding pattern.
  const double eps = std::max((double)margin, 4.5e-3);
  JoinLaunch jl;
  jl.n = m;
  jl.k_blocks = pitch / kBlockK;
  jl.part_index = part_index;
  jl.part_count = part_count;
  jl.threshold = (float)(floor_d - eps);
  jl.member_bitmap = w_jbits.p;
  jl.limit = o->limit > 0 ? o->limit : 0;
  if ((rc = w_pair_count.reserve(4))) return rc;
  if (jl.limit > 0) {
    if ((rc = w_run_cut.reserve(1))) return rc;
    if ((rc = w_ghist.reserve((size_t)kJoinHistBins))) return rc;
    jl.run_cut = w_run_cut.p;
    jl.ghist = w_ghist.p;
    jl.kind = o->require_same_kind ? w_kind_c.p : nullptr;
    jl.band = (float)(2.0 * eps);
    jl.hist_lo = jl.threshold;
    jl.hist_scale = (float)((double)kJoinHistBins / std::max(1e-6, 1.0 + eps - (double)jl.threshold));
  }
  const bool pair_kernel = cfg.force_path != 2;
  // CTA-pair kernel by default: the join runs for seconds under the board power cap, where a third less
  // L2->SM traffic per flop is worth 13 % (1.53 s vs 1.72-1.77 s at 1M x 1536); force_path = 2 keeps the
  // 1-CTA kernel
  const int64_t items = selfjoin_items(m, jl.k_blocks, part_index, part_count, pair_kernel);
  // chunks of ~50 ms so that a cancel request (checked between chunks, two in flight) is honoured quickly
  int64_t chunk = std::max<int64_t>(4096, ((int64_t)(pair_kernel ? 1 : 2) << 18) * 24 / std::max(1, jl.k_blocks));
  if (const char* e = getenv("OMR_SELFJOIN_CHUNK")) chunk = std::max<int64_t>(1, atoll(e));
  unsigned long long pcap = std::max<unsigned long long>(w_pairs.n, 1ull << 20);
  if (const char* e = getenv("OMR_SELFJOIN_PCAP")) pcap = std::max<unsigned long long>(1ull, strtoull(e, nullptr, 10));
  unsigned long long cnt = 0;
  for (int attempt = 0; attempt < 3; attempt++) {
    if ((rc = w_pairs.reserve((size_t)pcap))) return rc;
    jl.pairs = w_pairs.p;
    jl.cap = pcap;
    jl.pair_count = w_pair_count.p;
    OMR_CUDA_OK(cudaMemsetAsync(w_pair_count.p, 0, 32, st));
    if (jl.limit > 0) {
      const int32_t neg = kOrderedNegInf;
      jpin.p[4] = (unsigned long long)(uint32_t)neg;
      OMR_CUDA_OK(cudaMemcpyAsync(w_run_cut.p, jpin.p + 4, 4, cudaMemcpyHostToDevice, st));
      OMR_CUDA_OK(cudaMemsetAsync(w_ghist.p, 0, (size_t)kJoinHistBins * 4, st));
    }
    if (profiling) OMR_CUDA_OK(cudaEventRecord(jev0, st));
    int c = 0;
    for (int64_t first = 0; first < items; first += chunk, c++) {
      jl.chunk_first = first;
      jl.chunk_items = std::min(chunk, items - first);
      rc = pair_kernel ? launch_selfjoin_umma2(ta, jl, num_sms, st) : launch_selfjoin_umma(ta, tb, jl, num_sms, st);
      if (rc) return rc;
      st_launches++;
      OMR_CUDA_OK(cudaEventRecord(jchunk[c & 1], st));
      if (c >= 1) {
        OMR_CUDA_OK(cudaEventSynchronize(jchunk[(c - 1) & 1]));
        if (cancelled()) {
          OMR_CUDA_OK(cudaStreamSynchronize(st));
          return fail_msg(OMR_E_CANCELLED, "memory: self-join cancelled");
        }
      }
    }
    if (profiling) OMR_CUDA_OK(cudaEventRecord(jev1, st));
    OMR_CUDA_OK(cudaMemcpyAsync(jpin.p, w_pair_count.p, 8, cudaMemcpyDeviceToHost, st));
    OMR_CUDA_OK(cudaStreamSynchronize(st));
    cnt = jpin.p[0];
    if (profiling) {
      float ms = 0.0f;
      OMR_CUDA_OK(cudaEventElapsedTime(&ms, jev0, jev1));
      scorer_ms_sum += ms;
      scorer_timed++;
    }
    if (cancelled()) return fail_msg(OMR_E_CANCELLED, "memory: self-join cancelled");
    if (cnt <= pcap) break;
    if (attempt == 2) return fail_msg(OMR_E_CAPACITY, "memory: self-join candidate count kept growing");
    pcap = cnt + (cnt >> 3);   // the list was too short: rerun with room for what this pass counted
  }
  st_join_candidates += (int64_t)cnt;
  if (cnt == 0) return OMR_OK;

  // ---- exact similarity, row-pair rules, ORDER BY / LIMIT: all on the device
  if (!identity) {
    if ((rc = launch_pair_remap(w_pairs.p, cnt, w_idx.p, st))) return rc;
    st_launches++;
  }
  if ((rc = w_pair_sim.reserve((size_t)cnt))) return rc;
  if ((rc = w_keys.reserve((size_t)cnt * 2))) return rc;
  if ((rc = w_jstate.reserve(join_select_state_bytes()))) return rc;
  if ((rc = launch_pair_keys(view(), w_pairs.p, cnt, d_entity_key, d_kind, o->require_same_kind, floor_d, w_pair_sim.p,
                             w_keys.p, w_pair_count.p + 1, st)))
    return rc;
  st_launches++;
  OMR_CUDA_OK(cudaMemcpyAsync(jpin.p + 1, w_pair_count.p + 1, 8, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  const unsigned long long n_valid = jpin.p[1];
  if (n_valid == 0) return OMR_OK;
  const unsigned long long out_cap = jl.limit > 0 ? std::min<unsigned long long>(n_valid, (unsigned long long)jl.limit) : n_valid;
  if ((rc = w_out_pairs.reserve((size_t)out_cap))) return rc;
  if ((rc = w_out_sim.reserve((size_t)out_cap))) return rc;
  if ((rc = launch_pair_select(w_pairs.p, w_pair_sim.p, w_keys.p, cnt, jl.limit, w_jstate.p, out_cap, w_out_pairs.p,
                               w_out_sim.p, w_pair_count.p + 2, num_sms, st)))
    return rc;
  st_launches += jl.limit > 0 ? 17 : 1;
  out_pairs->resize((size_t)out_cap);
  out_sims->resize((size_t)out_cap);
  OMR_CUDA_OK(cudaMemcpyAsync(out_pairs->data(), w_out_pairs.p, (size_t)out_cap * sizeof(PairCand), cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaMemcpyAsync(out_sims->data(), w_out_sim.p, (size_t)out_cap * 8, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaMemcpyAsync(jpin.p + 2, w_pair_count.p + 2, 8, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  if (jpin.p[2] != out_cap)
    return fail_msg(OMR_E_CUDA, "memory: gpu: self-join selection kept %llu of %llu pairs", jpin.p[2], out_cap);
  // best first: similarity DESC, then (canonical row, dupe row), the order the selection used
  std::vector<size_t> order((size_t)out_cap);
  for (size_t i = 0; i < order.size(); i++) order[i] = i;
  std::sort(order.begin(), order.end(), [&](size_t x, size_t y) {
    if ((*out_sims)[x] != (*out_sims)[y]) return (*out_sims)[x] > (*out_sims)[y];
    if ((*out_pairs)[x].a != (*out_pairs)[y].a) return (*out_pairs)[x].a < (*out_pairs)[y].a;
    return (*out_pairs)[x].b < (*out_pairs)[y].b;
  });
  std::vector<PairCand> sp(order.size());
  std::vector<double> ss(order.size());
  for (size_t i = 0; i < order.size(); i++) {
    sp[i] = (*out_pairs)[order[i]];
    ss[i] = (*out_sims)[order[i]];
  }
  out_pairs->swap(sp);
  out_sims->swap(ss);
  return OMR_OK;
}

int omr_selfjoin(omr_store* s, const omr_selfjoin_opts* o, omr_pair* out, int64_t cap, int64_t* found) {
  if (!s || !o || !found || (cap > 0 && !out)) return fail_msg(OMR_E_INVALID, "memory: omr_selfjoin: null argument");
  if (!o->workspace_id || !o->workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  *found = 0;
  std::lock_guard<std::mutex> gate(s->join_gate);
  std::vector<PairCand> pairs;
  std::vector<double> sims;
  if (int rc = s->selfjoin(o, &pairs, &sims)) return rc;
  std::lock_guard<std::mutex> lk(s->mu);   // host identity maps
  const int64_t total = (int64_t)pairs.size();
  *found = total;
  for (int64_t i = 0; i < std::min(total, cap); i++) {
    omr_pair& p = out[i];
    key_to(s->entity_id[s->h_entity[pairs[i].a]], p.canonical_id);
    key_to(s->entity_id[s->h_entity[pairs[i].b]], p.dupe_id);
    p.similarity = sims[i];
    p.canonical_row = s->cfg.row_base + (int64_t)pairs[i].a * s->cfg.row_stride;
    p.dupe_row = s->cfg.row_base + (int64_t)pairs[i].b * s->cfg.row_stride;
    p.kind = s->h_kind[pairs[i].a];
  }
  return OMR_OK;
}

int omr_selfjoin_rows(omr_store* s, const omr_selfjoin_opts* o, int64_t* canonical_rows, int64_t* dupe_rows,
                      double* similarities, int64_t cap, int64_t* found) {
  if (!s || !o || !found || (cap > 0 && (!canonical_rows || !dupe_rows || !similarities)))
    return fail_msg(OMR_E_INVALID, "memory: omr_selfjoin_rows: null argument");
  if (!o->workspace_id || !o->workspace_id[0]) return fail_msg(OMR_E_WORKSPACE, "memory: workspace_id is required");
  *found = 0;
  std::lock_guard<std::mutex> gate(s->join_gate);
  std::vector<PairCand> pairs;
  std::vector<double> sims;
  if (int rc = s->selfjoin(o, &pairs, &sims)) return rc;
  const int64_t total = (int64_t)pairs.size();
  *found = total;
  for (int64_t i = 0; i < std::min(total, cap); i++) {
    canonical_rows[i] = s->cfg.row_base + (int64_t)pairs[i].a * s->cfg.row_stride;
    dupe_rows[i] = s->cfg.row_base + (int64_t)pairs[i].b * s->cfg.row_stride;
    similarities[i] = sims[i];
  }
  return OMR_OK;
}

int omr_row_entity(omr_store* s, int64_t row, uint8_t entity_id[16], uint32_t* kind) {
  if (!s || !entity_id) return fail_msg(OMR_E_INVALID, "memory: omr_row_entity: null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  const int64_t stride = s->cfg.row_stride > 0 ? s->cfg.row_stride : 1;
  const int64_t local = (row - s->cfg.row_base) / stride;
  if (row < s->cfg.row_base || (row - s->cfg.row_base) % stride != 0 || local >= s->n)
    return fail_msg(OMR_E_NOT_FOUND, "memory: not found");
  key_to(s->entity_id[s->h_entity[local]], entity_id);
  if (kind) *kind = s->h_kind[local];
  return OMR_OK;
}

// ---- PCA over resident rows (k_pca.cu) ---------------------------------------------------------

int omr_projection_rows(omr_store* s, const omr_filter* filter, int64_t* rows, int64_t cap, int64_t* found) {
  if (found) *found = 0;
  if (!s || !filter || (cap > 0 && !rows)) return fail_msg(OMR_E_INVALID, "memory: omr_projection_rows: null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  if (s->n == 0) return OMR_OK;
  DeviceFilter df;
  double unused;
  int rc;
  if ((rc = s->to_device_filter(*filter, &df, &unused))) return rc;
  if ((rc = s->w_filters.reserve(64))) return rc;
  if ((rc = s->w_bitmaps.reserve((size_t)s->words()))) return rc;
  OMR_CUDA_OK(cudaMemcpyAsync(s->w_filters.p, &df, sizeof df, cudaMemcpyHostToDevice, s->stream));
  if ((rc = launch_build_bitmaps(s->meta(), s->n, (int)s->words(), s->w_filters.p, 1, s->w_bitmaps.p, s->stream))) return rc;
  s->st_launches++;
  const size_t nw = (size_t)((s->n + 31) / 32);
  std::vector<uint32_t> bits(nw);
  OMR_CUDA_OK(cudaMemcpyAsync(bits.data(), s->w_bitmaps.p, nw * 4, cudaMemcpyDeviceToHost, s->stream));
  OMR_CUDA_OK(cudaStreamSynchronize(s->stream));
  // newest eligible row per entity: rows are appended in arrival order, so "newest" is the largest index
  std::vector<int64_t> best(s->entity_rows.size(), -1);
  for (int64_t r = 0; r < s->n; r++)
    if (bits[(size_t)(r >> 5)] >> (r & 31) & 1u) best[s->h_entity[r]] = r;
  std::vector<int64_t> picked;
  for (int64_t r : best)
    if (r >= 0) picked.push_back(r);
  std::sort(picked.begin(), picked.end());
  if (found) *found = (int64_t)picked.size();
  const int64_t w = std::min<int64_t>(cap, (int64_t)picked.size());
  for (int64_t i = 0; i < w; i++) rows[i] = picked[(size_t)i];
  return OMR_OK;
}

static int pca_stage_rows(omr_store* s, const int64_t* rows, int64_t n) {
  for (int64_t i = 0; i < n; i++)
    if (rows[i] < 0 || rows[i] >= s->n)
      return fail_msg(OMR_E_INVALID, "memory: projection row %lld outside the store (%lld rows)", (long long)rows[i],
                      (long long)s->n);
  int rc = s->w_out_rows.reserve((size_t)n);
  if (rc) return rc;
  OMR_CUDA_OK(cudaMemcpyAsync(s->w_out_rows.p, rows, (size_t)n * 8, cudaMemcpyHostToDevice, s->stream));
  return 0;
}

int omr_pca_moments(omr_store* s, const int64_t* rows, int64_t n, double* mean, double* cov) {
  if (!s || !rows || !mean || !cov) return fail_msg(OMR_E_INVALID, "memory: omr_pca_moments: null argument");
  if (n < 2) return fail_msg(OMR_E_INVALID, "memory: projection: PCA needs at least 2 rows, got %lld", (long long)n);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  int rc;
  if ((rc = pca_stage_rows(s, rows, n))) return rc;
  const int chunks = pca_mean_chunks(n);
  const int splits = pca_cov_splits(n, s->dim, s->num_sms);
  const size_t dd = (size_t)s->dim * s->dim;
  // layout of the fp64 scratch: mean | cov | column-sum partials | covariance partials
  const size_t need = (size_t)s->dim + dd + (size_t)chunks * s->dim + (size_t)splits * dd;
  if ((rc = s->w_exact.reserve(need))) return rc;
  double* d_mean = s->w_exact.p;
  double* d_cov = d_mean + s->dim;
  double* d_colsum = d_cov + dd;
  double* d_part = d_colsum + (size_t)chunks * s->dim;
  if ((rc = launch_pca_moments(s->view(), s->w_out_rows.p, n, d_colsum, chunks, d_mean, d_part, splits, d_cov, s->stream)))
    return rc;
  s->st_launches += 4;
  OMR_CUDA_OK(cudaMemcpyAsync(mean, d_mean, (size_t)s->dim * 8, cudaMemcpyDeviceToHost, s->stream));
  OMR_CUDA_OK(cudaMemcpyAsync(cov, d_cov, dd * 8, cudaMemcpyDeviceToHost, s->stream));
  OMR_CUDA_OK(cudaStreamSynchronize(s->stream));
  return OMR_OK;
}

int omr_pca_project(omr_store* s, const int64_t* rows, int64_t n, const double* basis, int32_t k, double* out) {
  if (!s || !rows || !basis || !out) return fail_msg(OMR_E_INVALID, "memory: omr_pca_project: null argument");
  if (n <= 0) return OMR_OK;
  if (k < 1 || k > s->dim) return fail_msg(OMR_E_INVALID, "memory: projection: k must be in 1..%d, got %d", s->dim, k);
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (int drc = s->drain()) return drc;
  int rc;
  if ((rc = pca_stage_rows(s, rows, n))) return rc;
  const size_t nb = (size_t)s->dim * k;
  if ((rc = s->w_exact.reserve(nb + (size_t)n * k))) return rc;
  double* d_basis = s->w_exact.p;
  double* d_out = d_basis + nb;
  OMR_CUDA_OK(cudaMemcpyAsync(d_basis, basis, nb * 8, cudaMemcpyHostToDevice, s->stream));
  if ((rc = launch_pca_project(s->view(), s->w_out_rows.p, n, d_basis, k, d_out, s->stream))) return rc;
  s->st_launches++;
  OMR_CUDA_OK(cudaMemcpyAsync(out, d_out, (size_t)n * k * 8, cudaMemcpyDeviceToHost, s->stream));
  OMR_CUDA_OK(cudaStreamSynchronize(s->stream));
  return OMR_OK;
}

// Diagnostics: runs only query preparation + one scorer and returns the raw packed keys
// ([nq][n_tiles * (256 / S) * 2] int32, host), so tests can check a scorer in isolation.
int omr_debug_candidate_keys(omr_store* s, const float* queries, int32_t nq, int32_t path, int32_t S,
                             int32_t* out_keys, int64_t out_len) {
  if (!s || !queries || !out_keys) return fail_msg(OMR_E_INVALID, "memory: omr_debug_candidate_keys: null argument");
  if (S != 32 && S != 64 && S != 128 && S != 256) return fail_msg(OMR_E_INVALID, "memory: bad sub-bucket %d", S);
  if (path != 1 && path != 2 && path != 4) return fail_msg(OMR_E_INVALID, "memory: path must be 1, 2 or 4");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (s->n == 0 || nq <= 0) return OMR_OK;
  cudaStream_t st = s->stream;
  int rc;
  const int n_tiles = (int)((s->n + kTileRows - 1) / kTileRows);
  const int64_t kpq = (int64_t)n_tiles * (kTileRows / S) * 2;
  if (out_len < (int64_t)nq * kpq) return fail_msg(OMR_E_CAPACITY, "memory: need %lld keys", (long long)((int64_t)nq * kpq));
  const int qtile = path == 4 ? 2 * kQueryTile : kQueryTile;
  const int qpad = path != 1 ? ((nq + qtile - 1) / qtile) * qtile : std::max(nq, 8);
  if ((rc = s->w_queries.reserve((size_t)nq * s->dim))) return rc;
  if ((rc = s->lanes[0].qn.reserve((size_t)qpad * s->pitch))) return rc;
  if ((rc = s->lanes[0].qnorm2.reserve((size_t)qpad))) return rc;
  if ((rc = s->lanes[0].cand.reserve((size_t)qpad * kpq))) return rc;
  if ((rc = s->ensure_live(st))) return rc;
  OMR_CUDA_OK(cudaMemcpyAsync(s->w_queries.p, queries, (size_t)nq * s->dim * 4, cudaMemcpyHostToDevice, st));
  if ((rc = launch_prep_queries(s->w_queries.p, nq, qpad, s->dim, s->pitch, s->lanes[0].qn.p, s->lanes[0].qnorm2.p, nullptr, st))) return rc;
  if (path == 1) {
    for (int q0 = 0; q0 < nq; q0 += 8) {
      const int c = std::min(8, nq - q0);
      if ((rc = launch_score_stream(s->view(), s->lanes[0].qn.p + (size_t)q0 * s->pitch, c, n_tiles, S, s->w_live.p,
                                    (int)s->words(), nullptr, s->lanes[0].cand.p + (size_t)q0 * kpq, s->num_sms, st)))
        return rc;
    }
  } else {
    if (s->lanes[0].tmap_q_ptr != s->lanes[0].qn.p || s->lanes[0].tmap_q_rows != s->lanes[0].qn.n / s->pitch) {
      if ((rc = s->make_tmap(&s->lanes[0].tmap_q, s->lanes[0].qn.p, s->lanes[0].qn.n / s->pitch, kQueryTile))) return rc;
      s->lanes[0].tmap_q_ptr = s->lanes[0].qn.p;
      s->lanes[0].tmap_q_rows = s->lanes[0].qn.n / s->pitch;
    }
    if (path == 4)
      rc = launch_score_umma2(s->lanes[0].tmap_q, s->tmap_rows_a, nq, qpad / (2 * kQueryTile), n_tiles, s->pitch / kBlockK, S,
                              s->w_live.p, (int)s->words(), nullptr, s->lanes[0].cand.p, s->num_sms, st);
    else
      rc = launch_score_umma(s->lanes[0].tmap_q, s->tmap_rows_b, nq, qpad / kQueryTile, n_tiles, s->pitch / kBlockK, S,
                             s->w_live.p, (int)s->words(), nullptr, s->lanes[0].cand.p, s->num_sms, st);
    if (rc) return rc;
  }
  OMR_CUDA_OK(cudaMemcpyAsync(out_keys, s->lanes[0].cand.p, (size_t)nq * kpq * 4, cudaMemcpyDeviceToHost, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  return OMR_OK;
}

int omr_get_stats(const omr_store* s, omr_stats* out) {
  if (!s || !out) return fail_msg(OMR_E_INVALID, "memory: null argument");
  out->kernel_launches = s->st_launches.load();
  out->searches = s->st_searches.load();
  out->queries = s->st_queries.load();
  out->rescued_rows = s->st_rescued.load();
  out->fallback_queries = s->st_fallback.load();
  out->last_path = s->last_path;
  out->last_sub_bucket = s->last_S;
  out->last_rescore_width = s->last_K2;
  out->num_sms = s->num_sms;
  out->host_calls = s->st_host_calls.load();
  out->host_batches = s->st_host_batches.load();
  out->scorer_ms_sum = s->scorer_ms_sum;
  out->scorer_timed = s->scorer_timed;
  out->listed_entries = s->st_listed.load();
  for (int i = 0; i < 3; i++) out->phase_ms_sum[i] = s->phase_ms_sum[i];
  out->scorer_kernel_ms_sum = s->kernel_ns_sum * 1e-6;
  out->scorer_kernel_timed = s->kernel_ns_n;
  out->scorer_gap_ms_sum = s->gap_ns_sum * 1e-6;
  out->post_span_ms_sum = s->post_ns_sum * 1e-6;
  out->post_end_after_scorer_ms_sum = s->post_lag_ns_sum * 1e-6;
  out->join_candidates = s->st_join_candidates.load();
  out->scorer_start_skew_ms_sum = s->kernel_start_skew_ns * 1e-6;
  out->scorer_end_skew_ms_sum = s->kernel_end_skew_ns * 1e-6;
  return OMR_OK;
}

// Diagnostics: device time (ms, CUDA events) of `iters` back-to-back launches of one scorer over the
// whole store with nq random-free (zero) queries already prepared; nothing downstream runs.
int omr_debug_time_scorer(omr_store* s, int32_t nq, int32_t path, int32_t S, int32_t iters, float* ms_out) {
  if (!s || !ms_out || iters < 1) return fail_msg(OMR_E_INVALID, "memory: omr_debug_time_scorer: bad argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  cudaStream_t st = s->stream;
  omr_store::Lane& L = s->lanes[0];
  int rc;
  const int n_tiles = (int)((s->n + kTileRows - 1) / kTileRows);
  const bool emit = S == 0;   // emission mode (top-10 thresholds): seed pass + seed kernel + scorer per iteration
  if (emit && path != 2 && path != 4) return fail_msg(OMR_E_INVALID, "memory: emission mode is a tensor-core path");
  const int64_t kpq = emit ? 0 : (int64_t)n_tiles * (kTileRows / S) * 2;
  const int qtile = path == 4 ? 2 * kQueryTile : kQueryTile;
  const int qpad = path != 1 ? ((nq + qtile - 1) / qtile) * qtile : std::max(nq, 8);
  if ((rc = s->w_queries.reserve((size_t)nq * s->dim))) return rc;
  if ((rc = L.qn.reserve((size_t)qpad * s->pitch))) return rc;
  if ((rc = L.qnorm2.reserve((size_t)qpad))) return rc;
  if (!emit && (rc = L.cand.reserve((size_t)qpad * kpq))) return rc;
  if ((rc = s->ensure_live(st))) return rc;
  EmitView ev;
  int seed_n = 0;
  int64_t seed_kpq = 0;
  if (emit && (rc = s->prepare_emit(L, qpad, 10, n_tiles, st, &ev, &seed_n, &seed_kpq))) return rc;
  // queries = the store's own first rows (real data: all-zero operands would flatter the power-limited clocks)
  if (nq > s->n) return fail_msg(OMR_E_INVALID, "memory: omr_debug_time_scorer needs nq <= rows");
  OMR_CUDA_OK(cudaMemcpy2DAsync(s->w_queries.p, (size_t)s->dim * 4, s->d_f32, (size_t)s->pitch * 4, (size_t)s->dim * 4, nq,
                                cudaMemcpyDeviceToDevice, st));
  if ((rc = launch_prep_queries(s->w_queries.p, nq, qpad, s->dim, s->pitch, L.qn.p, L.qnorm2.p, nullptr, st))) return rc;
  if (L.tmap_q_ptr != L.qn.p || L.tmap_q_rows != L.qn.n / s->pitch) {
    if ((rc = s->make_tmap(&L.tmap_q, L.qn.p, L.qn.n / s->pitch, kQueryTile))) return rc;
    L.tmap_q_ptr = L.qn.p;
    L.tmap_q_rows = L.qn.n / s->pitch;
  }
  for (int it = -1; it < iters; it++) {
    if (it == 0) OMR_CUDA_OK(cudaEventRecord(L.ev0, st));
    if (emit) {
      rc = launch_score_umma(L.tmap_q, s->tmap_rows_b, nq, qpad / kQueryTile, seed_n, s->pitch / kBlockK, 32, s->w_live.p,
                             (int)s->words(), nullptr, L.seed_keys.p, s->num_sms, st);
      if (!rc) rc = launch_emit_seed(L.seed_keys.p, seed_kpq, L.qnorm2.p, nq, qpad, ev, s->use_pdl, st);
      if (!rc)
        rc = path == 4 ? launch_score_umma2_emit(L.tmap_q, s->tmap_rows_a, nq, qpad / (2 * kQueryTile), n_tiles,
                                                 s->pitch / kBlockK, s->w_live.p, (int)s->words(), nullptr, ev, nullptr, s->num_sms, s->use_pdl, st)
                       : launch_score_umma_emit(L.tmap_q, s->tmap_rows_b, nq, qpad / kQueryTile, n_tiles, s->pitch / kBlockK,
                                                s->w_live.p, (int)s->words(), nullptr, ev, nullptr, s->num_sms, s->use_pdl, st);
    } else if (path == 4)
      rc = launch_score_umma2(L.tmap_q, s->tmap_rows_a, nq, qpad / (2 * kQueryTile), n_tiles, s->pitch / kBlockK, S,
                              s->w_live.p, (int)s->words(), nullptr, L.cand.p, s->num_sms, st);
    else if (path == 2)
      rc = launch_score_umma(L.tmap_q, s->tmap_rows_b, nq, qpad / kQueryTile, n_tiles, s->pitch / kBlockK, S,
                             s->w_live.p, (int)s->words(), nullptr, L.cand.p, s->num_sms, st);
    else
      rc = launch_score_stream(s->view(), L.qn.p, std::min(nq, 8), n_tiles, S, s->w_live.p, (int)s->words(), nullptr,
                               L.cand.p, s->num_sms, st);
    if (rc) return rc;
  }
  OMR_CUDA_OK(cudaEventRecord(L.ev1, st));
  OMR_CUDA_OK(cudaStreamSynchronize(st));
  float ms = 0;
  OMR_CUDA_OK(cudaEventElapsedTime(&ms, L.ev0, L.ev1));
  *ms_out = ms / iters;
  return OMR_OK;
}

int omr_debug_post_timing(omr_store* s, int64_t* out, int32_t nq) {
  if (!s || !out) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  cudaSetDevice(s->cfg.device);
  if (!s->lanes[0].timing.p || s->lanes[0].timing.n < (size_t)nq * 8) return fail_msg(OMR_E_INVALID, "memory: post timing was not recorded (set OMR_POST_TIMING=1)");
  OMR_CUDA_OK(cudaMemcpy(out, s->lanes[0].timing.p, (size_t)nq * 8 * sizeof(long long), cudaMemcpyDeviceToHost));
  return OMR_OK;
}

int omr_set_profiling(omr_store* s, int32_t on) {
  if (!s) return fail_msg(OMR_E_INVALID, "memory: null argument");
  std::lock_guard<std::mutex> lk(s->mu);
  s->profiling = on != 0;
  s->scorer_ms_sum = 0.0;
  s->scorer_timed = 0;
  for (double& v : s->phase_ms_sum) v = 0.0;
  s->kernel_ns_sum = s->kernel_start_skew_ns = s->kernel_end_skew_ns = 0.0;
  s->gap_ns_sum = s->post_ns_sum = s->post_lag_ns_sum = 0.0;
  s->last_main_end_ns = 0;
  s->kernel_ns_n = 0;
  return OMR_OK;
}

const char* omr_kind_name(const omr_store* s, uint32_t kind) {
  if (!s || kind >= s->kind_names.names.size()) return "";
  return s->kind_names.names[kind].c_str();
}

}  // extern "C"
