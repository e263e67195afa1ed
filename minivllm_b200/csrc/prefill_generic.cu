// prefill_generic.cu — varlen causal prefill attention, generic CUDA-core kernel.
//
// Covers what the tcgen05 kernel (prefill_tcgen05.cu) does not: fp32 tensors, head_dim other
// than 128, block sizes that are not a multiple of 64.  Same semantics as the flash call site
// of the reference (minivllm/models/layers/attention.py:396-433) and its eager twin (:467-504,
// :554-597): bottom-right aligned causal mask, GQA by h / (H/KVH), K/V either dense new-token
// tensors (cu_seqlens_k) or the paged caches through block_tables.
// One warp per (query row, head); lanes split head_dim; online softmax in fp32.
#include <math.h>

#include "common.cuh"

namespace mvb200 {

template <typename T>
__global__ void __launch_bounds__(128)
prefill_generic_kernel(const T* __restrict__ q, int64_t q_row_stride, const T* __restrict__ k,
                       const T* __restrict__ v, int64_t k_row_stride, int64_t v_row_stride,
                       const T* __restrict__ k_cache, const T* __restrict__ v_cache,
                       const int32_t* __restrict__ block_tables, int64_t bt_row_stride,
                       const int32_t* __restrict__ cu_q, const int32_t* __restrict__ cu_k,
                       T* __restrict__ out, int64_t out_row_stride, int num_seqs, int total_q,
                       int num_heads, int num_kv_heads, int head_dim, int block_size,
                       int64_t num_blocks, int max_blocks, float scale_log2) {
  constexpr int kMaxPerLane = 8;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * 4 + warp;
  const int h = blockIdx.y;
  if (row >= total_q) return;
  const int kvh = h / (num_heads / num_kv_heads);
  // owning sequence: cu_q[s] <= row < cu_q[s+1] (skips empty sequences)
  int s = 0;
  while (s < num_seqs - 1 && row >= cu_q[s + 1]) ++s;
  const int q0 = cu_q[s], lq = cu_q[s + 1] - q0;
  const int k0 = cu_k[s], lk = cu_k[s + 1] - k0;
  const int i = row - q0;
  T* orow = out + (int64_t)row * out_row_stride + (int64_t)h * head_dim;
  if (i >= lq) return;  // row beyond the last sequence
  const int limit = min(lk, i + (lk - lq) + 1);  // keys [0, limit)

  float qr[kMaxPerLane], acc[kMaxPerLane];
#pragma unroll
  for (int c = 0; c < kMaxPerLane; ++c) {
    const int d = lane + 32 * c;
    qr[c] = d < head_dim ? to_f32(q[(int64_t)row * q_row_stride + (int64_t)h * head_dim + d]) : 0.f;
    acc[c] = 0.f;
  }
  float m = -INFINITY, l = 0.f;
  for (int j = 0; j < limit; ++j) {
    const T* kp;
    const T* vp;
    bool ok = true;
    if (block_tables) {
      const int bid = block_tables[(int64_t)s * bt_row_stride + min(j / block_size, max_blocks - 1)];
      ok = bid >= 0 && bid < num_blocks;
      const int64_t base = ((int64_t)(ok ? bid : 0) * block_size + j % block_size) * num_kv_heads * head_dim +
                           (int64_t)kvh * head_dim;
      kp = k_cache + base;
      vp = v_cache + base;
    } else {
      kp = k + (int64_t)(k0 + j) * k_row_stride + (int64_t)kvh * head_dim;
      vp = v + (int64_t)(k0 + j) * v_row_stride + (int64_t)kvh * head_dim;
    }
    float dot = 0.f;
#pragma unroll
    for (int c = 0; c < kMaxPerLane; ++c) {
      const int d = lane + 32 * c;
      dot += (ok && d < head_dim) ? qr[c] * to_f32(kp[d]) : 0.f;
    }
    dot = warp_sum(dot) * scale_log2;
    const float m_new = fmaxf(m, dot);
    const float a = exp2f(m - m_new);
    const float pw = exp2f(dot - m_new);
    l = l * a + pw;
#pragma unroll
    for (int c = 0; c < kMaxPerLane; ++c) {
      const int d = lane + 32 * c;
      const float vv = (ok && d < head_dim) ? to_f32(vp[d]) : 0.f;
      acc[c] = acc[c] * a + pw * vv;
    }
    m = m_new;
  }
#pragma unroll
  for (int c = 0; c < kMaxPerLane; ++c) {
    const int d = lane + 32 * c;
    if (d < head_dim) orow[d] = from_f32<T>(l > 0.f ? acc[c] / l : 0.f);
  }
}

template <typename T>
static int launch_prefill_generic(const void* q, int64_t q_row_stride, const void* k, const void* v,
                                  int64_t k_row_stride, int64_t v_row_stride, const void* k_cache,
                                  const void* v_cache, const int32_t* block_tables,
                                  int64_t bt_row_stride, const int32_t* cu_q, const int32_t* cu_k,
                                  void* out, int64_t out_row_stride, int num_seqs, int total_q,
                                  int num_heads, int num_kv_heads, int head_dim, int block_size,
                                  int64_t num_blocks, int max_blocks, float scale,
                                  cudaStream_t stream) {
  dim3 grid((total_q + 3) / 4, num_heads);
  prefill_generic_kernel<T><<<grid, 128, 0, stream>>>(
      (const T*)q, q_row_stride, (const T*)k, (const T*)v, k_row_stride, v_row_stride,
      (const T*)k_cache, (const T*)v_cache, block_tables, bt_row_stride, cu_q, cu_k, (T*)out,
      out_row_stride, num_seqs, total_q, num_heads, num_kv_heads, head_dim, block_size, num_blocks,
      max_blocks, scale * kLog2e);
  return check_launch("prefill_generic_kernel");
}

int prefill_generic_dispatch(const void* q, int64_t q_row_stride, const void* k, const void* v,
                             int64_t k_row_stride, int64_t v_row_stride, const void* k_cache,
                             const void* v_cache, const int32_t* block_tables,
                             int64_t bt_row_stride, const int32_t* cu_q, const int32_t* cu_k,
                             void* out, int64_t out_row_stride, int num_seqs, int total_q,
                             int num_heads, int num_kv_heads, int head_dim, int block_size,
                             int64_t num_blocks, int max_blocks, float scale, int dtype,
                             cudaStream_t stream) {
  if (head_dim > 256) return fail(MVB200_EUNSUPPORTED, "prefill: head_dim %d > 256", head_dim);
  if (num_heads > 65535) return fail(MVB200_EUNSUPPORTED, "prefill: num_heads %d > 65535", num_heads);
  note_kernel("prefill_generic");
#define MVB_ARGS                                                                                   \
  q, q_row_stride, k, v, k_row_stride, v_row_stride, k_cache, v_cache, block_tables, bt_row_stride, \
      cu_q, cu_k, out, out_row_stride, num_seqs, total_q, num_heads, num_kv_heads, head_dim,        \
      block_size, num_blocks, max_blocks, scale, stream
  switch (dtype) {
    case MVB200_F32: return launch_prefill_generic<float>(MVB_ARGS);
    case MVB200_F16: return launch_prefill_generic<__half>(MVB_ARGS);
    case MVB200_BF16: return launch_prefill_generic<__nv_bfloat16>(MVB_ARGS);
  }
#undef MVB_ARGS
  return fail(MVB200_EINVAL, "prefill: unknown dtype %d", dtype);
}

}  // namespace mvb200
