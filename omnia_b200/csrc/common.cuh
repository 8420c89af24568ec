// Shared definitions for the recall engine's device code.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

namespace omr {

// Rows are scored in tiles of 256 (one UMMA N).  A tile is split into
// sub-buckets of S rows (S in {32, 64, 128, 256}); the scorers keep, per query
// and sub-bucket, the two best approximate scores as packed keys.
constexpr int kTileRows = 256;
constexpr int kQueryTile = 128;  // UMMA M: queries per accumulator tile
constexpr int kBlockK = 64;      // bf16 elements per 128-byte swizzle atom
constexpr int32_t kKeyNone = INT32_MIN;

// Row flag bits, device copy of the host metadata (include/omnia_recall.h).
constexpr uint8_t kRowForgotten = 1;
constexpr uint8_t kRowSuperseded = 2;
constexpr uint8_t kRowHasEmbedding = 4;
constexpr uint8_t kRowMutable = 8;
constexpr uint8_t kRowEntityRep = 16;
constexpr uint8_t kRowFree = 32;      // host mirror only: the slot of a hard-deleted row is on the free list
constexpr uint8_t kRowDeleted = 64;

// Packed candidate key: an order-preserving map of the fp32 score to int32
// with the low 8 bits replaced by the row's offset inside its sub-bucket.
// Larger key == better score; kKeyNone marks "nothing eligible".
__host__ __device__ __forceinline__ int32_t score_to_ordered(float s) {
#ifdef __CUDA_ARCH__
  int32_t b = __float_as_int(s);
#else
  int32_t b;
  memcpy(&b, &s, 4);
#endif
  return b ^ ((b >> 31) & 0x7fffffff);
}
__host__ __device__ __forceinline__ int32_t pack_key(float s, int col) {
  return (score_to_ordered(s) & ~0xFF) | col;
}
// Upper bound of the score a key can stand for (low byte saturated).
__host__ __device__ __forceinline__ float key_score_upper(int32_t key) {
  int32_t k = key | 0xFF;
  int32_t b = k ^ ((k >> 31) & 0x7fffffff);
#ifdef __CUDA_ARCH__
  return __int_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}

// Inverse of score_to_ordered.
__host__ __device__ __forceinline__ float ordered_to_score(int32_t k) {
  int32_t b = k ^ ((k >> 31) & 0x7fffffff);
#ifdef __CUDA_ARCH__
  return __int_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}
// The same order on unsigned integers (what the emission lists store): 0 is below every score.
__host__ __device__ __forceinline__ uint32_t score_to_biased(float s) { return (uint32_t)score_to_ordered(s) ^ 0x80000000u; }
__host__ __device__ __forceinline__ float biased_to_score(uint32_t u) { return ordered_to_score((int32_t)(u ^ 0x80000000u)); }
constexpr int32_t kOrderedNegInf = (int32_t)0x807FFFFF;  // score_to_ordered(-inf): every score clears it
constexpr int32_t kOrderedPosInf = (int32_t)0x7F800000;  // score_to_ordered(+inf): no finite score clears it

__device__ __forceinline__ void top2_insert(int32_t& m1, int32_t& m2, int32_t k) {
  m2 = max(m2, min(m1, k));
  m1 = max(m1, k);
}

// Device-side view of one store shard.  All pointers are device memory owned by
// the Store; `pitch` is the row pitch in ELEMENTS for both matrices.
struct ShardView {
  const __nv_bfloat16* rows_bf16;  // [cap][pitch]  m / |m| rounded to bf16
  const float* rows_f32;           // [cap][pitch]  the vectors exactly as ingested
  const double* norm2;             // [cap]         sum m_i^2 in fp64
  int64_t n;                       // rows in use
  int dim;
  int pitch;
};

// Row metadata, struct-of-arrays (reference row a8: the columns the recall
// WHERE clause reads, internal/memory/store.go:250-258).
struct RowMetaView {
  const uint32_t* workspace;
  const uint32_t* user;   // 0 = NULL
  const uint32_t* agent;  // 0 = NULL
  const uint16_t* kind;
  const uint16_t* purpose;
  const float* confidence;
  const uint8_t* flags;
  const int64_t* valid_until_us;  // INT64_MAX = NULL
  const double* norm2;            // 0 (or non-finite) = the vector has no cosine
};

constexpr int kMaxFilterSet = 16;
struct DeviceFilter {
  uint32_t workspace;
  uint32_t user;
  int32_t user_mode;
  uint32_t agent;
  int32_t agent_mode;
  int32_t n_kinds;
  int32_t n_purposes;
  uint16_t kinds[kMaxFilterSet];
  uint16_t purposes[kMaxFilterSet];
  float min_confidence;
  int64_t now_us;
};

#define OMR_CUDA_OK(expr)                                                              \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) return ::omr::fail_cuda(_e, #expr, __FILE__, __LINE__);     \
  } while (0)

int fail_cuda(cudaError_t e, const char* what, const char* file, int line);
int fail_msg(int code, const char* fmt, ...);

}  // namespace omr
