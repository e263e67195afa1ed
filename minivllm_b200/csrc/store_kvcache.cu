// store_kvcache.cu — coalesced scatter of new K/V rows into the paged cache.
//
// Replaces StandardAttentionBackend.store_kv_cache (reference
// minivllm/models/layers/attention_backend.py:196-266) and the Triton store_kvcache_kernel
// (:52-78).  HBM-bound byte mover: K and V go in one launch, every thread moves one 16-byte
// vector of the K row and the matching vector of the V row, consecutive threads touch
// consecutive vectors of the same token row (fully coalesced 128-bit accesses on both sides).
// Algorithmic traffic per token: row_bytes * 2 (K,V) * 2 (read+write) + 4 (slot).
#include "common.cuh"

namespace mvb200 {

template <typename Vec>
__global__ void __launch_bounds__(256)
store_kvcache_kernel(const char* __restrict__ key, const char* __restrict__ value,
                     int64_t key_row_bytes_stride, int64_t value_row_bytes_stride,
                     char* __restrict__ k_cache, char* __restrict__ v_cache,
                     const int32_t* __restrict__ slot_mapping, int64_t total_vecs,
                     int32_t vecs_per_row, int64_t num_slots) {
  pdl_wait();  // key/value/slot_mapping come from the kernels before; the cache may still be read by them
  pdl_launch_dependents();
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t row_bytes = (int64_t)vecs_per_row * sizeof(Vec);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_vecs; i += stride) {
    const int64_t tok = i / vecs_per_row;
    const int32_t vec = (int32_t)(i - tok * vecs_per_row);
    const int64_t slot = __ldg(slot_mapping + tok);
    if (slot < 0 || slot >= num_slots) continue;  // -1 = padding row (attention_backend.py:70,239)
    const Vec kv = *reinterpret_cast<const Vec*>(key + tok * key_row_bytes_stride + (int64_t)vec * sizeof(Vec));
    const Vec vv = *reinterpret_cast<const Vec*>(value + tok * value_row_bytes_stride + (int64_t)vec * sizeof(Vec));
    *reinterpret_cast<Vec*>(k_cache + slot * row_bytes + (int64_t)vec * sizeof(Vec)) = kv;
    *reinterpret_cast<Vec*>(v_cache + slot * row_bytes + (int64_t)vec * sizeof(Vec)) = vv;
  }
}

template <typename Vec>
static int launch_store(const void* key, const void* value, int64_t ks_bytes, int64_t vs_bytes,
                        void* k_cache, void* v_cache, const int32_t* slots, int32_t n,
                        int64_t row_bytes, int64_t num_slots, cudaStream_t stream) {
  const int32_t vpr = (int32_t)(row_bytes / sizeof(Vec));
  const int64_t total = (int64_t)n * vpr;
  const int threads = 256;
  int64_t blocks = (total + threads - 1) / threads;
  const int64_t cap = 148LL * 32;  // grid-stride beyond 32 CTAs per SM
  if (blocks > cap) blocks = cap;
  cudaError_t e = launch_pdl(store_kvcache_kernel<Vec>, dim3((unsigned)blocks), dim3(threads), 0, stream,
                             (const char*)key, (const char*)value, ks_bytes, vs_bytes,
                             (char*)k_cache, (char*)v_cache, slots, total, vpr, num_slots);
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
