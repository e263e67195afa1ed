/*
 * mvb200.h — C ABI of the B200-native attention hot path for mini-vLLM.
 *
 * One shared library (libmvb200.so), plain pointers and sizes, no torch types.
 * Every launcher is stateless and re-entrant: it enqueues kernels on `stream`
 * (a cudaStream_t passed as void*), never allocates device memory, never
 * synchronises the host, and is therefore CUDA-Graph capturable.  Return value:
 * 0 on success, a non-zero MVB200_E* code otherwise; mvb200_last_error() gives
 * the message for the calling thread.
 *
 * The reference (jianzhnie/mini-vllm) has no native code; each entry point
 * replaces a Python/Triton/flash-attn call site of the reference, cited as
 * file:line relative to the reference repository root.
 *
 * Layouts (all row-major, element strides unless said otherwise):
 *   key/value/q/out : [N, heads, D], the two inner dims dense, row stride given
 *   k_cache/v_cache : [NB, BS, KVH, D] contiguous            (inference_executor.py:209-221)
 *   slot_mapping    : int32 [N], slot = block*BS + offset, <0 = skip (inference_executor.py:503-510,568-572)
 *   block_tables    : int32 [S, MB] padded with -1            (inference_executor.py:619-688)
 *   context_lens    : int32 [B], includes the current token   (inference_executor.py:566)
 *   cu_seqlens_q/k  : int32 [S+1]                             (inference_executor.py:519-524)
 */
#ifndef MVB200_H_
#define MVB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define MVB200_API __attribute__((visibility("default")))
#else
#define MVB200_API
#endif

#define MVB200_VERSION 130

/* element types */
#define MVB200_F32 0
#define MVB200_F16 1
#define MVB200_BF16 2

/* error codes */
#define MVB200_OK 0
#define MVB200_EINVAL 1      /* bad argument (shape, stride, alignment, null pointer) */
#define MVB200_EUNSUPPORTED 2 /* valid request this build has no kernel for */
#define MVB200_ECUDA 3       /* CUDA runtime / driver error at launch */

MVB200_API int mvb200_version(void);
MVB200_API const char* mvb200_last_error(void);

/* Name of the kernel family the last successful call on this thread dispatched to
 * ("decode_tma_mma", "decode_generic", "prefill_tcgen05", "prefill_generic", "store_vec16", ...). */
MVB200_API const char* mvb200_last_kernel(void);

/* Comma-separated list of the distinct kernel families this process has dispatched to so far, written
 * to buf (NUL-terminated, truncated to cap); returns their count.  Diagnostic: lets a host engine that
 * swapped the layer in prove which kernels its run actually launched. */
MVB200_API int mvb200_kernels_seen(char* buf, int cap);

/*
 * store_kvcache — scatter new K/V rows into the paged cache.
 * Replaces StandardAttentionBackend.store_kv_cache (models/layers/attention_backend.py:196-266)
 * and the Triton store_kvcache_kernel (models/layers/attention_backend.py:52-78, launch :225-234).
 *
 * For every i < num_tokens with 0 <= slot_mapping[i] < num_slots:
 *   k_cache[slot_mapping[i], 0:row_elems] = key[i*key_row_stride + 0:row_elems]   (same for value)
 * Bits are moved, never converted.  row_elems = KVH*D.
 */
MVB200_API int mvb200_store_kvcache(const void* key, const void* value,
                         int64_t key_row_stride, int64_t value_row_stride,
                         void* k_cache, void* v_cache,
                         const int32_t* slot_mapping,
                         int32_t num_tokens, int32_t row_elems, int64_t num_slots,
                         int32_t dtype, void* stream);

/*
 * paged decode attention — one query token per sequence over its paged KV.
 * Replaces flash_attn_with_kvcache(...) at models/layers/attention.py:439-447 and the eager
 * PageAttention.__call__ (models/layers/page_attention.py:35-133).
 *
 * out[b,h,:] = softmax(scale * q[b,h,:] . K_b[0:ctx_b]^T) V_b[0:ctx_b],  kv head = h / (H/KVH).
 * Rows with context_lens[b] <= 0 produce zeros.  Block-table entries that are negative inside
 * the context contribute a zero key/value row (the eager reference's behaviour).
 *
 * num_splits >= 1 partitions every sequence's KV range into that many work units (the grid is
 * persistent and independent of the data, as CUDA Graph replay needs).  The CTA that finishes the last
 * split of a (sequence, kv head) combines the partials: one kernel, no separate reduction launch.
 *
 * workspace: mvb200_decode_workspace_bytes(B, H, D, num_splits) bytes, 16-byte aligned, size a multiple
 * of 16, prepared ONCE with mvb200_decode_workspace_init (which zeroes the arrival tickets and the work
 * counter kept in the LAST bytes of the buffer and stamps it; the kernel re-arms them itself, so the same
 * buffer serves every later call, any batch size or split count it is large enough for — always pass the
 * same workspace_bytes).  One workspace per stream: two launches in flight must not share one.
 * Required when num_splits > 1.  With num_splits == 1 it is optional: without it the units are dealt
 * round-robin instead of being drawn from the device-side counter (ragged batches balance worse).
 */
MVB200_API size_t mvb200_decode_workspace_bytes(int32_t batch, int32_t num_heads, int32_t head_dim,
                                     int32_t num_splits);
MVB200_API int mvb200_decode_workspace_init(void* workspace, size_t workspace_bytes, void* stream);

/* Split count the library recommends for a batch (>=1). max_context bounds the longest sequence. */
MVB200_API int32_t mvb200_decode_suggest_splits(int32_t batch, int32_t num_kv_heads, int32_t max_context,
                                     int32_t sm_count);

MVB200_API int mvb200_paged_decode(const void* q, int64_t q_row_stride,
                        const void* k_cache, const void* v_cache,
                        const int32_t* block_tables, int64_t bt_row_stride,
                        const int32_t* context_lens,
                        void* out, int64_t out_row_stride,
                        void* workspace, size_t workspace_bytes,
                        int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim,
                        int32_t block_size, int64_t num_blocks, int32_t max_blocks_per_seq,
                        int32_t num_splits, float scale, int32_t dtype, void* stream);

/* Flags of mvb200_paged_decode_ex. */
/* The caller vouches that the kernel launched immediately before this one on the stream writes neither
 * block_tables / context_lens nor any K/V row other than the NEWEST token of each sequence (position
 * context_lens[b] - 1) — true for the layer's decode path, whose predecessor is the store_kvcache of the
 * step's new tokens (attention.py:237-243).  Under programmatic dependent launch the kernel then reads the
 * metadata and starts streaming the older K/V tiles while that predecessor is still draining — so context_lens
 * and block_tables must be final BEFORE the predecessor starts, not merely before this call. */
#define MVB200_DECODE_KV_PREFIX_STABLE 1u

/* The step's new tokens, to be stored by the decode kernel itself (fused store_kvcache, saves one launch per
 * layer): row b of key/value [B, KVH, D] goes to slot_mapping[b] when 0 <= slot < num_blocks * block_size,
 * bits moved unchanged, exactly as mvb200_store_kvcache would.  slot_mapping[b] must be the slot of position
 * context_lens[b] - 1 of sequence b (what _prepare_decode_input produces, inference_executor.py:568-572) or
 * negative (padding row).  Shapes the fused path does not cover fall back to a separate store launch. */
typedef struct mvb200_decode_store {
  const void* key;
  const void* value;
  int64_t key_row_stride;   /* elements between token rows */
  int64_t value_row_stride;
  const int32_t* slot_mapping;
} mvb200_decode_store;

/* mvb200_paged_decode with flags and an optional fused store (flags = 0, store = NULL: the call above). */
MVB200_API int mvb200_paged_decode_ex(const void* q, int64_t q_row_stride,
                           const void* k_cache, const void* v_cache,
                           const int32_t* block_tables, int64_t bt_row_stride,
                           const int32_t* context_lens,
                           void* out, int64_t out_row_stride,
                           void* workspace, size_t workspace_bytes,
                           int32_t batch, int32_t num_heads, int32_t num_kv_heads, int32_t head_dim,
                           int32_t block_size, int64_t num_blocks, int32_t max_blocks_per_seq,
                           int32_t num_splits, float scale, int32_t dtype, uint32_t flags,
                           const mvb200_decode_store* store, void* stream);

/*
 * varlen causal prefill attention.
 * Replaces flash_attn_varlen_func(...) at models/layers/attention.py:422-433 and the eager
 * Attention._fallback_attention prefill branch (models/layers/attention.py:467-504).
 *
 * Sequence s owns query rows cu_seqlens_q[s]..cu_seqlens_q[s+1]-1 (Lq of them) and keys
 * 0..Lk-1, Lk = cu_seqlens_k[s+1]-cu_seqlens_k[s].  Query i attends keys 0..(Lk-Lq)+i
 * (bottom-right aligned causal mask, attention.py:584-587).
 *
 * Key/value source:
 *   block_tables != NULL : key j of sequence s is k_cache[block_tables[s, j/BS], j%BS]
 *                          (the prefix-cache path, attention.py:399-400); k/v are ignored.
 *   block_tables == NULL : key j is k[(cu_seqlens_k[s]+j)*k_row_stride] (dense new-token K/V).
 */
MVB200_API int mvb200_prefill_varlen(const void* q, int64_t q_row_stride,
                          const void* k, const void* v, int64_t k_row_stride, int64_t v_row_stride,
                          int64_t num_kv_rows,
                          const void* k_cache, const void* v_cache,
                          const int32_t* block_tables, int64_t bt_row_stride,
                          const int32_t* cu_seqlens_q, const int32_t* cu_seqlens_k,
                          void* out, int64_t out_row_stride,
                          int32_t num_seqs, int32_t total_q, int32_t max_seqlen_q, int32_t max_seqlen_k,
                          int32_t num_heads, int32_t num_kv_heads, int32_t head_dim,
                          int32_t block_size, int64_t num_blocks, int32_t max_blocks_per_seq,
                          float scale, int32_t dtype, void* stream);

/*
 * qk_norm_rope_store — caller-side fusion in front of the attention layer.
 * Replaces the sequence in QwenAttention.forward, models/qwen_base.py:181-186:
 *   q_norm / k_norm  (RMSNorm over head_dim, models/layers/layernorm.py:81-93; Qwen3 only)
 *   rotary_emb       (models/layers/rotary_embedding.py:166-193, apply_rotary_emb :78-84)
 * and the store Attention.forward begins with (models/layers/attention.py:228-243,
 * models/layers/attention_backend.py:196-266).
 *
 * For every token i and head h:   x = q[i,h,:] (or k[i,h,:])
 *   if weights given: x = T( float(x) * (1/sqrt(mean(float(x)^2) + eps)) * float(w) )
 *   c, s = cos_cache[pos], sin_cache[pos]   pos = clamp(positions[i], 0, max_positions-1), tables fp32 [max_positions, D/2]
 *   out[0:D/2] = T( x1*c - x2*s ),  out[D/2:D] = T( x2*c + x1*s )      x1, x2 = halves of float(x)
 * q_out / k_out receive the result (they may alias q / k: in place; k_out may be NULL), and for
 * 0 <= slot_mapping[i] < num_slots the rotated k row and the untouched v row are written to
 * k_cache / v_cache row slot_mapping[i] ([num_slots, KVH*D]).  slot_mapping NULL: no store.
 * q_norm_weight / k_norm_weight: [D] in the model dtype, both NULL for models without q/k norm.
 */
MVB200_API int mvb200_qk_norm_rope_store(const void* q, int64_t q_row_stride,
                          const void* k, int64_t k_row_stride,
                          const void* v, int64_t v_row_stride,
                          void* q_out, int64_t q_out_row_stride,
                          void* k_out, int64_t k_out_row_stride,
                          const int64_t* positions,
                          const float* cos_cache, const float* sin_cache, int64_t max_positions,
                          const void* q_norm_weight, const void* k_norm_weight, float eps,
                          const int32_t* slot_mapping, void* k_cache, void* v_cache, int64_t num_slots,
                          int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads,
                          int32_t head_dim, int32_t dtype, void* stream);

/*
 * decode_advance — advance the persistent decode metadata by one token, on the device.
 * Replaces, per step, InferenceExecutor._prepare_decode_input (engine/inference_executor.py:551-617),
 * _prepare_block_tables_decode (:651-688) and the graph-buffer refill of _execute_with_cuda_graph
 * (:751-768).  For every row b < batch with context_lens[b] > 0 (len = context_lens[b] + 1):
 *   if appends[b] >= 0: block_tables[b, (len-1)/block_size] = appends[b]     (the block handed out by
 *                                                     BlockManager.may_append when len % BS == 1)
 *   input_ids[b] = next_tokens[b];  positions[b] = len-1;  context_lens[b] = len
 *   slot_mapping[b] = block_tables[b, (len-1)/block_size]*block_size + (len-1)%block_size   (:568-572)
 *                     (-1 when that entry is negative, >= num_blocks or beyond max_blocks)
 * Rows with context_lens[b] <= 0 are graph-bucket padding: input_ids = positions = 0, slot = -1 (:757-760).
 * appends[b] is reset to -1 for every row.  All tensors are the persistent (graph) buffers; no host sync.
 */
MVB200_API int mvb200_decode_advance(const int64_t* next_tokens, int32_t* appends, int64_t* input_ids,
                          int64_t* positions, int32_t* slot_mapping, int32_t* context_lens,
                          int32_t* block_tables, int64_t bt_row_stride, int32_t batch,
                          int32_t max_blocks, int32_t block_size, int64_t num_blocks, void* stream);

/*
 * greedy_argmax — the temperature == 0 branch of the reference sampler (sampling/sampler.py:104-106;
 * GreedySampler :118-130): out_index[r] = index_offset + argmax_j logits[r, j], ties to the lowest j,
 * NaN counts as the maximum (torch.argmax).  out_value (optional, fp32) receives the maximum itself so
 * vocabulary-sharded ranks can combine their winners.  workspace: mvb200_greedy_argmax_workspace_bytes.
 */
MVB200_API size_t mvb200_greedy_argmax_workspace_bytes(int32_t rows, int32_t vocab);
MVB200_API int mvb200_greedy_argmax(const void* logits, int64_t row_stride, int32_t rows, int32_t vocab,
                          int32_t dtype, void* workspace, size_t workspace_bytes,
                          int64_t* out_index, float* out_value, int64_t index_offset, void* stream);

/*
 * greedy_argmax_static — mvb200_greedy_argmax for callers that have no scratch to give and may not allocate.
 * No counterpart in the reference (its sampler runs on torch tensors and torch's allocator); it exists for
 * hosts of this library that cannot obtain device memory at call time: a serving process that handed every
 * spare byte of HBM to the KV cache when it started, a binding from another language (cgo / JNI / a bare
 * ctypes shim) that carries no device allocator of its own, or a step loop that cannot afford the
 * device-wide stall a cudaMalloc brings.  Same arguments and same result as mvb200_greedy_argmax minus the
 * workspace pair; logits, out_index and out_value are still the caller's.
 *
 * The per-chunk partial winners go to an array compiled into the library (MVB200_GREEDY_ARGMAX_STATIC_BYTES,
 * 256-byte aligned).  Each GPU that loads libmvb200.so holds that much memory until the process exits.  The
 * call never runs cudaMalloc, cudaFree or any allocating routine.
 *   limit     : when the vocabulary is cut into more than one chunk per row, the rows x chunks partial keys
 *               (mvb200_greedy_argmax_workspace_bytes(rows, vocab)) must fit MVB200_GREEDY_ARGMAX_STATIC_BYTES,
 *               otherwise MVB200_EINVAL with both sizes in the message.  The present chunking never asks for
 *               more than 590 keys = 4720 bytes (rows = 295), and from 296 rows on uses one chunk and no
 *               scratch, so every rows x vocab is accepted today; the check keeps a later change of the
 *               chunking from writing past the array.
 *   first call: the first call on each device looks up the array's device address (cudaGetSymbolAddress,
 *               with the caller's current device — make the device that owns `stream` current).  Make that
 *               first call outside any stream capture.
 *   sharing   : all calls on a device use the same array and it keeps whatever the last call wrote.  Nothing
 *               in it has to start at zero — the first pass rewrites every key the second pass reads — but
 *               two calls that overlap on the GPU (different streams, or graph branches running side by side)
 *               race on it and can return each other's winners.  Keep calls on one device ordered.
 */
#define MVB200_GREEDY_ARGMAX_STATIC_BYTES 8192
MVB200_API int mvb200_greedy_argmax_static(const void* logits, int64_t row_stride, int32_t rows, int32_t vocab,
                          int32_t dtype, int64_t* out_index, float* out_value, int64_t index_offset,
                          void* stream);

/*
 * Peer all-reduce fused with residual add + RMSNorm.
 * Replaces, for decode-size messages, dist.all_reduce in RowParallelLinear.forward
 * (models/layers/linear.py:314-319) and the RMSNorm.forward(x, residual) that follows it
 * (models/layers/layernorm.py:81-105):
 *   s = T( sum over ranks (rank order, fp32) of x_rank )            -- the all-reduce result, model dtype
 *   if residual: r = float(s) + float(residual); residual = T(r)    else r = float(s)
 *   y = weight ? T( r * (1/sqrt(mean(r^2) + eps)) * float(weight) ) : T(r)
 * Every rank calls it with the same rows/hidden in the same order; every rank gets identical y/residual.
 *
 * Buffers (per rank, cudaMalloc'ed by mvb200_peer_alloc so they can be exported over CUDA IPC):
 *   data : 2 * (world + 2) * slot_bytes  (per call parity: world partial slots, one Y and one R region)
 *   flags: mvb200_peer_flag_bytes()    (zero-initialised; [source rank][CTA] epochs)
 *   epoch: mvb200_peer_epoch_bytes()   (zero-initialised, local only)
 * data_ptrs[i] / flag_ptrs[i] are rank i's buffers as mapped into THIS process (own pointers at
 * index `rank`; peers through mvb200_peer_export -> mvb200_peer_open).  rows*hidden*sizeof(T) <= slot_bytes,
 * hidden*sizeof(T) a multiple of 16 and <= 16384.  No host sync, no allocation: CUDA-Graph capturable.
 * algo 1 = one-shot (every rank pushes its whole partial to every peer; latency-bound sizes),
 * algo 2 = two-shot (row r is reduced and normalised by rank (r % cap) % world only, which then pushes the
 * finished row to everybody: (world-1)/world of the message per direction).  With algo 2 and
 * gather_residual == 0 the residual stays row-sharded: a rank updates only the rows it owns.
 * All ranks must pass the same algo / gather_residual / max_ctas for a given call.
 * Grid = min(rows, max_ctas or mvb200_peer_max_ctas()) CTAs, CTA c owning rows c, c+grid, ...; max_ctas
 * (0 = library default) must be the same on every rank and for every call on a communicator.
 * A peer that never arrives trips a device-side timeout (trap) after ~20 s instead of hanging the GPU.
 */
MVB200_API int mvb200_peer_flag_bytes(void);
MVB200_API int mvb200_peer_epoch_bytes(void);
MVB200_API int mvb200_peer_max_ctas(void);
MVB200_API int mvb200_peer_alloc(size_t bytes, void** ptr);          /* cudaMalloc + zero */
MVB200_API int mvb200_peer_free(void* ptr);
MVB200_API int mvb200_peer_export(void* ptr, void* handle64);        /* cudaIpcGetMemHandle, 64 bytes */
MVB200_API int mvb200_peer_open(const void* handle64, void** ptr);   /* cudaIpcOpenMemHandle */
MVB200_API int mvb200_peer_close(void* ptr);
MVB200_API int mvb200_peer_allreduce_add_rms_norm(void* const* data_ptrs, void* const* flag_ptrs, void* epoch,
                          int64_t slot_bytes, int32_t rank, int32_t world,
                          const void* x, int64_t x_row_stride,
                          void* residual, int64_t residual_row_stride,
                          const void* weight, void* y, int64_t y_row_stride,
                          int32_t rows, int32_t hidden, float eps, int32_t dtype, int32_t max_ctas,
                          int32_t algo, int32_t gather_residual, void* stream);

/*
 * The same operator with the NVSwitch doing the sum and the broadcast (NVLS / multimem): two-shot ownership,
 * multimem.ld_reduce of the owner's rows from a multicast mapping, multimem.st of the finished rows.
 *   multicast_base / local_base: the multicast mapping and this rank's own mapping of ONE symmetric buffer of
 *   6 * slot_bytes ([2 parities][X | Y | R][slot_bytes]) that every rank allocated identically (e.g. with
 *   torch.distributed._symmetric_memory: handle.multicast_ptr / tensor.data_ptr()).
 *   flag_ptrs / epoch: the flag and epoch buffers of mvb200_peer_allreduce_add_rms_norm (shared with it).
 * world >= 2; every rank makes the same call.  Replaces linear.py:314-319 + layernorm.py:81-105 like the call above.
 */
MVB200_API int mvb200_nvls_allreduce_add_rms_norm(void* multicast_base, void* local_base, int64_t slot_bytes,
                                       void* const* flag_ptrs, void* epoch, int32_t rank, int32_t world,
                                       const void* x, int64_t x_row_stride, void* residual,
                                       int64_t residual_row_stride, const void* weight, void* y,
                                       int64_t y_row_stride, int32_t rows, int32_t hidden, float eps, int32_t dtype,
                                       int32_t max_ctas, int32_t gather_residual, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MVB200_H_ */
