// prefill_api.cu — C entry point of the varlen causal prefill attention and kernel dispatch.
#include <stdlib.h>

#include "common.cuh"
#include "prefill_args.h"

namespace mvb200 {


int prefill_generic_dispatch(const void* q, int64_t q_row_stride, const void* k, const void* v,
                             int64_t k_row_stride, int64_t v_row_stride, const void* k_cache,
                             const void* v_cache, const int32_t* block_tables,
                             int64_t bt_row_stride, const int32_t* cu_q, const int32_t* cu_k,
                             void* out, int64_t out_row_stride, int num_seqs, int total_q,
                             int num_heads, int num_kv_heads, int head_dim, int block_size,
                             int64_t num_blocks, int max_blocks, float scale, int dtype,
                             cudaStream_t stream);

// return -1 when the request is outside what the kernel covers
int prefill_fa2q_dispatch(const PrefillArgs& a, cudaStream_t stream);     // persistent, two query tiles per CTA
int prefill_dbs_dispatch(const PrefillArgs& a, cudaStream_t stream);      // + 64-key steps with double-buffered S
#ifdef MVB200_TUNING
// first-generation kernel (one query tile per CTA): A/B tool of the tuning build only
// (python minivllm_b200/build.py --trace), not part of libmvb200.so
int prefill_tcgen05_dispatch(const PrefillArgs& a, cudaStream_t stream);
#endif

}  // namespace mvb200

extern "C" int mvb200_prefill_varlen(
    const void* q, int64_t q_row_stride, const void* k, const void* v, int64_t k_row_stride,
    int64_t v_row_stride, int64_t num_kv_rows, const void* k_cache, const void* v_cache,
    const int32_t* block_tables, int64_t bt_row_stride, const int32_t* cu_seqlens_q,
    const int32_t* cu_seqlens_k, void* out, int64_t out_row_stride, int32_t num_seqs,
    int32_t total_q, int32_t max_seqlen_q, int32_t max_seqlen_k, int32_t num_heads,
    int32_t num_kv_heads, int32_t head_dim, int32_t block_size, int64_t num_blocks,
    int32_t max_blocks_per_seq, float scale, int32_t dtype, void* stream) {
  using namespace mvb200;
  if (num_seqs < 0 || total_q < 0 || num_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0)
    return fail(MVB200_EINVAL, "prefill: bad sizes");
  if (num_heads % num_kv_heads)
    return fail(MVB200_EINVAL, "prefill: num_heads %d not a multiple of num_kv_heads %d", num_heads,
                num_kv_heads);
  if (num_seqs == 0 || total_q == 0) return MVB200_OK;
  if (!q || !out || !cu_seqlens_q || !cu_seqlens_k)
    return fail(MVB200_EINVAL, "prefill: null pointer");
  if (block_tables) {
    if (!k_cache || !v_cache || block_size <= 0 || num_blocks <= 0 || max_blocks_per_seq <= 0)
      return fail(MVB200_EINVAL, "prefill: block_tables given without a usable cache");
  } else if (!k || !v) {
    return fail(MVB200_EINVAL, "prefill: neither block_tables nor dense k/v given");
  }
  if (dtype != MVB200_F32 && dtype != MVB200_F16 && dtype != MVB200_BF16)
    return fail(MVB200_EINVAL, "prefill: unknown dtype %d", dtype);
  cudaStream_t s = (cudaStream_t)stream;

  PrefillArgs a{q, q_row_stride, k, v, k_row_stride, v_row_stride, num_kv_rows, k_cache, v_cache,
                block_tables, bt_row_stride, cu_seqlens_q, cu_seqlens_k, out, out_row_stride,
                num_seqs, total_q, max_seqlen_q, max_seqlen_k, num_heads, num_kv_heads, head_dim,
                block_size, num_blocks, max_blocks_per_seq, scale, dtype};
#ifdef MVB200_TUNING
  const bool force_generic = getenv("MVB200_PREFILL_GENERIC") != nullptr;
  const bool force_v1 = getenv("MVB200_PREFILL_V1") != nullptr;
  if (!force_generic) {
    const char* which = getenv("MVB200_PREFILL_KERNEL");  // tuning build: fa2q | dbs (default)
    int rc = (!force_v1 && !(which && which[0] == 'f')) ? prefill_dbs_dispatch(a, s) : -1;
    if (rc >= 0) return rc;
    rc = force_v1 ? -1 : prefill_fa2q_dispatch(a, s);
    if (rc >= 0) return rc;
    rc = prefill_tcgen05_dispatch(a, s);
    if (rc >= 0) return rc;
  }
#else
  {
    // double-buffered-S kernel first (355k vs 389k SM cycles on the cfg3 prefill shape, profiles/r02_prefill_*_run14);
    // the 128-key-tile kernel covers what it declines
    int rc = prefill_dbs_dispatch(a, s);
    if (rc >= 0) return rc;
    rc = prefill_fa2q_dispatch(a, s);
    if (rc >= 0) return rc;
  }
#endif
  return prefill_generic_dispatch(q, q_row_stride, k, v, k_row_stride, v_row_stride, k_cache,
                                  v_cache, block_tables, bt_row_stride, cu_seqlens_q, cu_seqlens_k,
                                  out, out_row_stride, num_seqs, total_q, num_heads, num_kv_heads,
                                  head_dim, block_size, num_blocks, max_blocks_per_seq, scale, dtype,
                                  s);
}
