// prefill_args.h — argument pack shared by the prefill entry point and its kernels.
#pragma once
#include <stdint.h>

namespace mvb200 {

struct PrefillArgs {
  const void* q;
  int64_t q_row_stride;
  const void* k;
  const void* v;
  int64_t k_row_stride, v_row_stride;
  int64_t num_kv_rows;
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;
  int64_t bt_row_stride;
  const int32_t* cu_q;
  const int32_t* cu_k;
  void* out;
  int64_t out_row_stride;
  int32_t num_seqs, total_q, max_seqlen_q, max_seqlen_k;
  int32_t num_heads, num_kv_heads, head_dim, block_size;
  int64_t num_blocks;
  int32_t max_blocks;
  float scale;
  int32_t dtype;
};

}  // namespace mvb200
