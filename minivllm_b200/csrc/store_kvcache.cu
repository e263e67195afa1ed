// store_kvcache.cu — coalesced scatter of new K/V rows into the paged cache.
//
// Replaces StandardAttentionBackend.store_kv_cache (reference
// minivllm/models/layers/attention_backend.py:196-266) and the Triton store_kvcache_kernel
// (:52-78).  HBM-bound byte mover: K and V go in one launch; a CTA owns a tile of token rows, consecutive
// threads touch consecutive 16-byte vectors of the same row (fully coalesced 128-bit accesses on both
// sides) and every thread keeps 128 B of loads in flight before its first store.
// Algorithmic traffic per token: row_bytes * 2 (K,V) * 2 (read+write) + 4 (slot).
#include "common.cuh"

namespace mvb200 {

// read-only path load of a vector that is read exactly once
template <typename Vec>
__device__ __forceinline__ Vec ld_stream(const char* p) {
  return __ldg(reinterpret_cast<const Vec*>(p));
}

constexpr int kStoreThreads = 256;
constexpr int kStoreMaxRows = 256;  // token rows per CTA (their slots are staged in shared memory)
constexpr int kStoreUnroll = 4;     // K and V vectors a thread keeps in flight: 2 * 4 * 16 B = 128 B

// One CTA moves a tile of `rows_per_cta` consecutive token rows.  The slots of the tile are read once
// (coalesced) into shared memory, then the threads sweep the tile's vectors: consecutive threads ->
// consecutive 16-byte vectors of the same row on both sides, kStoreUnroll K loads + kStoreUnroll V loads
// issued before the first store so that every thread has 128 B in flight (a per-vector slot lookup and a
// 64-bit divide per vector, as in the first version, left short rows latency-bound).  Row / column of a
// vector are advanced incrementally (step_rows = 256 / vecs_per_row, step_cols = 256 % vecs_per_row).
template <typename Vec>
__global__ void __launch_bounds__(kStoreThreads)
store_kvcache_kernel(const char* __restrict__ key, const char* __restrict__ value,
                     int64_t key_row_bytes_stride, int64_t value_row_bytes_stride,
                     char* __restrict__ k_cache, char* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int32_t num_tokens,
                     int32_t vecs_per_row, int32_t rows_per_cta, int32_t step_rows, int32_t step_cols,
                     int64_t num_slots) {
  __shared__ int64_t s_slot[kStoreMaxRows];
  pdl_wait();  // key/value/slot_mapping come from the kernels before; the cache may still be read by them
  const int32_t row0 = blockIdx.x * rows_per_cta;
  const int32_t rows = min(rows_per_cta, num_tokens - row0);
  for (int32_t r = threadIdx.x; r < rows; r += kStoreThreads) {
    const int64_t slot = __ldg(slot_mapping + row0 + r);
    s_slot[r] = (slot < 0 || slot >= num_slots) ? -1 : slot;  // -1 = padding row (attention_backend.py:70,239)
  }
  __syncthreads();
  pdl_launch_dependents();
  const int64_t row_bytes = (int64_t)vecs_per_row * sizeof(Vec);
  const int32_t total = rows * vecs_per_row;
  int32_t row = threadIdx.x / vecs_per_row;
  int32_t col = threadIdx.x - row * vecs_per_row;
  for (int32_t base = threadIdx.x; base < total; base += kStoreThreads * kStoreUnroll) {
    Vec kreg[kStoreUnroll], vreg[kStoreUnroll];
    int64_t dst[kStoreUnroll];
#pragma unroll
    for (int u = 0; u < kStoreUnroll; ++u) {
      dst[u] = -1;
      if (base + u * kStoreThreads < total) {
        const int64_t slot = s_slot[row];
        if (slot >= 0) {
          const int64_t off = (int64_t)col * sizeof(Vec);
          kreg[u] = ld_stream<Vec>(key + (int64_t)(row0 + row) * key_row_bytes_stride + off);
          vreg[u] = ld_stream<Vec>(value + (int64_t)(row0 + row) * value_row_bytes_stride + off);
          dst[u] = slot * row_bytes + off;
        }
      }
      row += step_rows;
      col += step_cols;
      if (col >= vecs_per_row) {
        col -= vecs_per_row;
        ++row;
      }
    }
#pragma unroll
    for (int u = 0; u < kStoreUnroll; ++u) {
      if (dst[u] >= 0) {
        *reinterpret_cast<Vec*>(k_cache + dst[u]) = kreg[u];
        *reinterpret_cast<Vec*>(v_cache + dst[u]) = vreg[u];
      }
    }
  }
}

template <typename Vec>
static int launch_store(const void* key, const void* value, int64_t ks_bytes, int64_t vs_bytes,
                        void* k_cache, void* v_cache, const int32_t* slots, int32_t n,
                        int64_t row_bytes, int64_t num_slots, cudaStream_t stream) {
  if (row_bytes / (int64_t)sizeof(Vec) > (1 << 22))
    return fail(MVB200_EINVAL, "store_kvcache: row of %lld bytes is too long", (long long)row_bytes);
  const int32_t vpr = (int32_t)(row_bytes / sizeof(Vec));
  // tile: about 16 KiB of K (and of V) per CTA, fewer rows when that would leave SMs without a CTA
  int32_t rows = (int32_t)((16384 + row_bytes - 1) / row_bytes);
  const int32_t spread = (n + 4 * device_sm_count() - 1) / (4 * device_sm_count());
  if (rows > spread) rows = spread;
  if (rows > kStoreMaxRows) rows = kStoreMaxRows;
  if (rows < 1) rows = 1;
  const int64_t blocks = ((int64_t)n + rows - 1) / rows;
  cudaError_t e = launch_pdl(store_kvcache_kernel<Vec>, dim3((unsigned)blocks), dim3(kStoreThreads), 0, stream,
                             (const char*)key, (const char*)value, ks_bytes, vs_bytes,
                             (char*)k_cache, (char*)v_cache, slots, n, vpr, rows,
                             (int32_t)(kStoreThreads / vpr), (int32_t)(kStoreThreads % vpr), num_slots);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "store_kvcache: %s", cudaGetErrorString(e));
  return check_launch("store_kvcache");
}

}  // namespace mvb200

extern "C" int mvb200_store_kvcache(const void* key, const void* value, int64_t key_row_stride,
                                    int64_t value_row_stride, void* k_cache, void* v_cache,
                                    const int32_t* slot_mapping, int32_t num_tokens,
                                    int32_t row_elems, int64_t num_slots, int32_t dtype,
                                    void* stream) {
  using namespace mvb200;
  if (num_tokens < 0 || row_elems <= 0 || num_slots < 0)
    return fail(MVB200_EINVAL, "store_kvcache: bad sizes (N=%d row=%d slots=%lld)", num_tokens,
                row_elems, (long long)num_slots);
  if (num_tokens == 0 || num_slots == 0) return MVB200_OK;
  if (!key || !value || !k_cache || !v_cache || !slot_mapping)
    return fail(MVB200_EINVAL, "store_kvcache: null pointer");
  const int eb = dtype == MVB200_F32 ? 4 : ((dtype == MVB200_F16 || dtype == MVB200_BF16) ? 2 : 0);
  if (!eb) return fail(MVB200_EINVAL, "store_kvcache: unknown dtype %d", dtype);
  if (key_row_stride < row_elems || value_row_stride < row_elems)
    return fail(MVB200_EINVAL, "store_kvcache: row stride smaller than the row");
  const int64_t row_bytes = (int64_t)row_elems * eb;
  const int64_t ksb = key_row_stride * eb, vsb = value_row_stride * eb;
  cudaStream_t s = (cudaStream_t)stream;
  // widest vector every row start and the row length are aligned to
  const uintptr_t mix = (uintptr_t)key | (uintptr_t)value | (uintptr_t)k_cache |
                        (uintptr_t)v_cache | (uintptr_t)ksb | (uintptr_t)vsb | (uintptr_t)row_bytes;
  if ((mix & 15) == 0) {
    note_kernel("store_vec16");
    return launch_store<uint4>(key, value, ksb, vsb, k_cache, v_cache, slot_mapping, num_tokens,
                               row_bytes, num_slots, s);
  } else if ((mix & 7) == 0) {
    note_kernel("store_vec8");
    return launch_store<uint2>(key, value, ksb, vsb, k_cache, v_cache, slot_mapping, num_tokens,
                               row_bytes, num_slots, s);
  } else if ((mix & 3) == 0) {
    note_kernel("store_vec4");
    return launch_store<uint32_t>(key, value, ksb, vsb, k_cache, v_cache, slot_mapping,
                                  num_tokens, row_bytes, num_slots, s);
  }
  note_kernel("store_vec2");
  return launch_store<uint16_t>(key, value, ksb, vsb, k_cache, v_cache, slot_mapping, num_tokens,
                                row_bytes, num_slots, s);
}
