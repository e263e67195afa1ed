// decode_meta.cu — device-side decode bookkeeping: the per-step attention metadata update and the
// greedy (argmax) token pick, both capturable as nodes of the decode CUDA graph.
//
// decode_advance_kernel replaces the per-step host work of the reference's
// InferenceExecutor._prepare_decode_input / _prepare_block_tables_decode
// (minivllm/engine/inference_executor.py:551-617, :651-688) and the graph-buffer refill of
// _execute_with_cuda_graph (:751-768): instead of rebuilding every row of every tensor on the host
// and copying five tensors per step, the persistent device tensors are advanced by one token:
//     input_id = last sampled token, position = len-1, context_len = len,
//     slot = block_table[(len-1)/BS]*BS + (len-1)%BS                      (:563-572)
// with at most one block-table entry appended per sequence (the block the BlockManager hands out when
// len % BS == 1, block_manager.py `may_append`).  Integer work, bit-exact against the host version.
//
// greedy_argmax_kernel is the temperature == 0 branch of the reference sampler
// (minivllm/sampling/sampler.py:104-106, GreedySampler :118-130): argmax over the vocabulary, ties to
// the lowest index, NaN counts as the maximum (torch.argmax semantics).  HBM-bound row reduction:
// one CTA per (row, vocabulary chunk) with 16-byte loads, a second tiny pass when chunks > 1.
#include <math.h>

#include <mutex>

#include "common.cuh"

namespace mvb200 {

__global__ void __launch_bounds__(128)
decode_advance_kernel(const int64_t* __restrict__ next_tokens, int32_t* __restrict__ appends,
                      int64_t* __restrict__ input_ids, int64_t* __restrict__ positions,
                      int32_t* __restrict__ slot_mapping, int32_t* __restrict__ context_lens,
                      int32_t* __restrict__ block_tables, int64_t bt_row_stride, int32_t batch,
                      int32_t max_blocks, int32_t block_size, int64_t num_blocks) {
  pdl_wait();  // next_tokens is written by the argmax node of the previous replay
  pdl_launch_dependents();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= batch) return;
  const int32_t len = context_lens[b];
  const int32_t app = appends[b];
  appends[b] = -1;  // consumed: the host only writes rows that get a new block
  if (len <= 0) {   // padding row of the graph bucket (inference_executor.py:757-760)
    input_ids[b] = 0;
    positions[b] = 0;
    slot_mapping[b] = -1;
    return;
  }
  const int32_t new_len = len + 1;
  const int32_t last = new_len - 1;  // index of the token fed this step
  const int32_t col = last / block_size;
  int32_t* row = block_tables + (int64_t)b * bt_row_stride;
  int32_t slot = -1;
  if (col < max_blocks) {
    if (app >= 0) row[col] = app;
    const int32_t blk = row[col];
    if (blk >= 0 && blk < num_blocks) slot = blk * block_size + last % block_size;
  }
  input_ids[b] = next_tokens[b];
  positions[b] = last;
  slot_mapping[b] = slot;
  context_lens[b] = new_len;
}

// ---- greedy argmax --------------------------------------------------------------------------------
// Order-preserving key: larger float -> larger key, NaN above +inf; low word = ~index so that the
// 64-bit max picks the lowest index among equal values.
MVB_DEV unsigned long long argmax_key(float v, uint32_t idx) {
  if (v == 0.f) v = 0.f;  // -0.0 == +0.0: the tie goes to the lower index, as in torch.argmax
  uint32_t u = __float_as_uint(v);
  uint32_t ord;
  if (v != v) ord = 0xFFFFFFFFu;
  else ord = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)ord << 32) | (unsigned long long)(0xFFFFFFFFu - idx);
}
MVB_DEV float key_value(unsigned long long key) {
  const uint32_t ord = (uint32_t)(key >> 32);
  if (ord == 0xFFFFFFFFu) return __uint_as_float(0x7FC00000u);
  const uint32_t u = (ord & 0x80000000u) ? (ord & 0x7FFFFFFFu) : ~ord;
  return __uint_as_float(u);
}
MVB_DEV unsigned long long umax64(unsigned long long a, unsigned long long b) { return a > b ? a : b; }

template <typename T>
struct Vec16;  // 16-byte vector of T and its unpack
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  MVB_DEV static void unpack(const uint4& r, float (&f)[4]) {
    f[0] = __uint_as_float(r.x); f[1] = __uint_as_float(r.y);
    f[2] = __uint_as_float(r.z); f[3] = __uint_as_float(r.w);
  }
};
template <>
struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  MVB_DEV static void unpack(const uint4& r, float (&f)[8]) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  }
};
template <>
struct Vec16<__half> {
  static constexpr int N = 8;
  MVB_DEV static void unpack(const uint4& r, float (&f)[8]) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
      const float2 v = __half22float2(h);
      f[2 * i] = v.x;
      f[2 * i + 1] = v.y;
    }
  }
};

constexpr int kArgmaxThreads = 512;

MVB_DEV unsigned long long block_max_key(unsigned long long key) {
  __shared__ unsigned long long warp_best[kArgmaxThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) key = umax64(key, __shfl_xor_sync(0xffffffffu, key, o));
  if ((threadIdx.x & 31) == 0) warp_best[threadIdx.x >> 5] = key;
  __syncthreads();
  if (threadIdx.x < 32) {
    key = threadIdx.x < kArgmaxThreads / 32 ? warp_best[threadIdx.x] : 0ull;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) key = umax64(key, __shfl_xor_sync(0xffffffffu, key, o));
  }
  return key;  // valid in thread 0
}

// grid (chunks, rows).  chunks == 1: writes the result; otherwise the per-chunk key goes to `partial`.
template <typename T>
__global__ void __launch_bounds__(kArgmaxThreads)
greedy_argmax_kernel(const T* __restrict__ logits, int64_t row_stride, int32_t vocab,
                     int32_t chunk_elems, unsigned long long* __restrict__ partial,
                     int64_t* __restrict__ out_index, float* __restrict__ out_value,
                     int64_t index_offset) {
  pdl_wait();
  pdl_launch_dependents();
  constexpr int N = Vec16<T>::N;
  const int row = blockIdx.y;
  const T* base = logits + (int64_t)row * row_stride;
  const int32_t lo = blockIdx.x * chunk_elems;
  const int32_t hi = min(vocab, lo + chunk_elems);
  // Per thread: running (value, index) with a strict '>' — indices are visited in increasing order, so the
  // first maximum wins — plus the lowest NaN index; ~6 ALU ops per element instead of a 64-bit key per
  // element (the kernel was ALU-bound at 1.7 TB/s).  Keys are built once per thread for the reductions.
  constexpr uint32_t kNone = 0xFFFFFFFFu;
  float best_v = -INFINITY;
  uint32_t best_i = kNone, nan_i = kNone, first_i = kNone;
  auto visit = [&](float v, uint32_t idx) {
    const bool gt = v > best_v;  // false for NaN
    best_v = gt ? v : best_v;
    best_i = gt ? idx : best_i;
    nan_i = (v != v) ? min(nan_i, idx) : nan_i;
  };
  const bool aligned = ((reinterpret_cast<uintptr_t>(base) & 15) == 0) && (lo % N == 0);
  int32_t vec_end = lo;
  if (aligned) {
    const int32_t nvec = (hi - lo) / N;
    vec_end = lo + nvec * N;
    if ((int32_t)threadIdx.x < nvec) first_i = (uint32_t)(lo + threadIdx.x * N);
    const uint4* p = reinterpret_cast<const uint4*>(base + lo);
    constexpr int U = 4;  // 64 bytes in flight per thread
    for (int32_t i0 = threadIdx.x; i0 < nvec; i0 += U * kArgmaxThreads) {
      uint4 r[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int32_t i = i0 + u * kArgmaxThreads;
        r[u] = i < nvec ? __ldg(p + i) : make_uint4(0, 0, 0, 0);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int32_t i = i0 + u * kArgmaxThreads;
        if (i < nvec) {
          float f[N];
          Vec16<T>::unpack(r[u], f);
#pragma unroll
          for (int j = 0; j < N; ++j) visit(f[j], (uint32_t)(lo + i * N + j));
        }
      }
    }
  }
  if (first_i == kNone && vec_end + (int32_t)threadIdx.x < hi) first_i = (uint32_t)(vec_end + threadIdx.x);
  for (int32_t i = vec_end + threadIdx.x; i < hi; i += kArgmaxThreads) visit(to_f32(base[i]), (uint32_t)i);
  // thread result -> key: a NaN beats everything; a thread that saw only -inf offers its first element
  unsigned long long best = 0ull;
  if (nan_i != kNone) best = argmax_key(__uint_as_float(0x7FC00000u), nan_i);
  else if (best_i != kNone) best = argmax_key(best_v, best_i);
  else if (first_i != kNone) best = argmax_key(-INFINITY, first_i);
  best = block_max_key(best);
  if (threadIdx.x == 0) {
    if (gridDim.x == 1) {
      out_index[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)best) + index_offset;
      if (out_value) out_value[row] = key_value(best);
    } else {
      partial[(int64_t)row * gridDim.x + blockIdx.x] = best;
    }
  }
}

__global__ void __launch_bounds__(32)
greedy_argmax_finish_kernel(const unsigned long long* __restrict__ partial, int32_t chunks,
                            int64_t* __restrict__ out_index, float* __restrict__ out_value,
                            int64_t index_offset) {
  pdl_wait();
  pdl_launch_dependents();
  const int row = blockIdx.x;
  unsigned long long best = 0ull;
  for (int i = threadIdx.x; i < chunks; i += 32) best = umax64(best, partial[(int64_t)row * chunks + i]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = umax64(best, __shfl_xor_sync(0xffffffffu, best, o));
  if (threadIdx.x == 0) {
    out_index[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)best) + index_offset;
    if (out_value) out_value[row] = key_value(best);
  }
}

static int32_t argmax_chunks(int32_t rows, int32_t vocab) {
  // enough CTAs to cover the 148 SMs twice, chunks of at least 8192 elements, at most 64 per row
  int32_t c = (2 * 148 + rows - 1) / rows;
  const int32_t by_size = (vocab + 8191) / 8192;
  if (c > by_size) c = by_size;
  if (c > 64) c = 64;
  return c < 1 ? 1 : c;
}

template <typename T>
static int launch_argmax(const void* logits, int64_t row_stride, int32_t rows, int32_t vocab,
                         void* workspace, int64_t* out_index, float* out_value, int64_t index_offset,
                         cudaStream_t s) {
  const int32_t chunks = argmax_chunks(rows, vocab);
  int32_t chunk_elems = (vocab + chunks - 1) / chunks;
  chunk_elems = (chunk_elems + 7) / 8 * 8;  // keeps every chunk start 16-byte aligned
  cudaError_t e = launch_pdl(greedy_argmax_kernel<T>, dim3(chunks, rows), dim3(kArgmaxThreads), 0, s,
                             (const T*)logits, row_stride, vocab, chunk_elems,
                             (unsigned long long*)workspace, out_index, out_value, index_offset);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "greedy_argmax: %s", cudaGetErrorString(e));
  if (chunks > 1) {
    e = launch_pdl(greedy_argmax_finish_kernel, dim3(rows), dim3(32), 0, s,
                   (const unsigned long long*)workspace, chunks, out_index, out_value, index_offset);
    if (e != cudaSuccess) return fail(MVB200_ECUDA, "greedy_argmax_finish: %s", cudaGetErrorString(e));
  }
  return check_launch("greedy_argmax");
}

}  // namespace mvb200

extern "C" int mvb200_decode_advance(const int64_t* next_tokens, int32_t* appends, int64_t* input_ids,
                                     int64_t* positions, int32_t* slot_mapping, int32_t* context_lens,
                                     int32_t* block_tables, int64_t bt_row_stride, int32_t batch,
                                     int32_t max_blocks, int32_t block_size, int64_t num_blocks,
                                     void* stream) {
  using namespace mvb200;
  if (batch < 0 || max_blocks < 0 || block_size <= 0 || num_blocks < 0 || bt_row_stride < max_blocks)
    return fail(MVB200_EINVAL, "decode_advance: bad sizes (B=%d MB=%d BS=%d NB=%lld stride=%lld)", batch,
                max_blocks, block_size, (long long)num_blocks, (long long)bt_row_stride);
  if (batch == 0) return MVB200_OK;
  if (!next_tokens || !appends || !input_ids || !positions || !slot_mapping || !context_lens ||
      (!block_tables && max_blocks > 0))
    return fail(MVB200_EINVAL, "decode_advance: null pointer");
  note_kernel("decode_advance");
  const int threads = 128;
  cudaError_t e = launch_pdl(decode_advance_kernel, dim3((batch + threads - 1) / threads), dim3(threads), 0,
                             (cudaStream_t)stream, next_tokens, appends, input_ids, positions,
                             slot_mapping, context_lens, block_tables, bt_row_stride, batch, max_blocks,
                             block_size, num_blocks);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "decode_advance: %s", cudaGetErrorString(e));
  return check_launch("decode_advance");
}

extern "C" size_t mvb200_greedy_argmax_workspace_bytes(int32_t rows, int32_t vocab) {
  if (rows <= 0 || vocab <= 0) return 0;
  return (size_t)rows * (size_t)mvb200::argmax_chunks(rows, vocab) * sizeof(unsigned long long);
}

extern "C" int mvb200_greedy_argmax(const void* logits, int64_t row_stride, int32_t rows, int32_t vocab,
                                    int32_t dtype, void* workspace, size_t workspace_bytes,
                                    int64_t* out_index, float* out_value, int64_t index_offset,
                                    void* stream) {
  using namespace mvb200;
  if (rows < 0 || vocab <= 0 || row_stride < vocab)
    return fail(MVB200_EINVAL, "greedy_argmax: bad sizes (rows=%d vocab=%d stride=%lld)", rows, vocab,
                (long long)row_stride);
  if (rows == 0) return MVB200_OK;
  if (!logits || !out_index) return fail(MVB200_EINVAL, "greedy_argmax: null pointer");
  if (argmax_chunks(rows, vocab) > 1 &&
      (!workspace || workspace_bytes < mvb200_greedy_argmax_workspace_bytes(rows, vocab)))
    return fail(MVB200_EINVAL, "greedy_argmax: workspace too small (%zu < %zu)", workspace_bytes,
                mvb200_greedy_argmax_workspace_bytes(rows, vocab));
  cudaStream_t s = (cudaStream_t)stream;
  note_kernel("greedy_argmax");
  switch (dtype) {
    case MVB200_BF16:
      return launch_argmax<__nv_bfloat16>(logits, row_stride, rows, vocab, workspace, out_index, out_value, index_offset, s);
    case MVB200_F16:
      return launch_argmax<__half>(logits, row_stride, rows, vocab, workspace, out_index, out_value, index_offset, s);
    case MVB200_F32:
      return launch_argmax<float>(logits, row_stride, rows, vocab, workspace, out_index, out_value, index_offset, s);
    default:
      return fail(MVB200_EINVAL, "greedy_argmax: unknown dtype %d", dtype);
  }
}

// ---- greedy argmax without a caller workspace -------------------------------------------------------
// Some hosts of this library cannot hand us scratch memory and must not have us allocate any: a serving
// process that gave all spare HBM to the KV cache at start-up, a host-language binding (cgo / JNI / ctypes shim)
// that owns no device allocator, or a step loop where cudaMalloc's implicit device synchronisation is not
// acceptable.  For them the per-chunk winners of the two-pass argmax live in the array below, which is part of
// the library image: every GPU that loads libmvb200.so carries these 8 KiB until the process exits.
//
// Host code must never pass the NAME of this array to a kernel — on the host the name is only a handle of the
// runtime, not a device address.  static_argmax_scratch() asks the runtime for the real address, once per
// device, with that device current.
#define MVB200_STATIC_ARGMAX_KEYS (MVB200_GREEDY_ARGMAX_STATIC_BYTES / sizeof(unsigned long long))
static __device__ __align__(256) unsigned long long g_static_argmax_partial[MVB200_STATIC_ARGMAX_KEYS];

namespace mvb200 {

constexpr int kMaxDevices = 64;
static std::mutex g_static_argmax_mu;
static unsigned long long* g_static_argmax_ptr[kMaxDevices];  // device address of the array, per device ordinal

// Device address of g_static_argmax_partial on the CURRENT device (resolved on the first call per device).
static int static_argmax_scratch(unsigned long long** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "greedy_argmax_static: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(MVB200_EUNSUPPORTED, "greedy_argmax_static: device ordinal %d beyond %d", dev, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_static_argmax_mu);
  if (!g_static_argmax_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_argmax_partial);
    if (e != cudaSuccess || !p) {
      (void)cudaGetLastError();
      return fail(MVB200_ECUDA, "greedy_argmax_static: cudaGetSymbolAddress on device %d: %s", dev,
                  cudaGetErrorString(e));
    }
    g_static_argmax_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = g_static_argmax_ptr[dev];
  return MVB200_OK;
}

}  // namespace mvb200

extern "C" int mvb200_greedy_argmax_static(const void* logits, int64_t row_stride, int32_t rows, int32_t vocab,
                                           int32_t dtype, int64_t* out_index, float* out_value,
                                           int64_t index_offset, void* stream) {
  using namespace mvb200;
  if (rows < 0 || vocab <= 0 || row_stride < vocab)
    return fail(MVB200_EINVAL, "greedy_argmax_static: bad sizes (rows=%d vocab=%d stride=%lld)", rows, vocab,
                (long long)row_stride);
  if (rows == 0) return MVB200_OK;
  if (!logits || !out_index) return fail(MVB200_EINVAL, "greedy_argmax_static: null pointer");
  if (dtype != MVB200_BF16 && dtype != MVB200_F16 && dtype != MVB200_F32)
    return fail(MVB200_EINVAL, "greedy_argmax_static: unknown dtype %d", dtype);
  // a single chunk per row writes the result directly and touches no scratch at all
  const size_t need = argmax_chunks(rows, vocab) > 1 ? mvb200_greedy_argmax_workspace_bytes(rows, vocab) : 0;
  if (need > (size_t)MVB200_GREEDY_ARGMAX_STATIC_BYTES)
    return fail(MVB200_EINVAL,
                "greedy_argmax_static: rows=%d x vocab=%d needs %zu bytes of scratch, the built-in array holds %d; "
                "use mvb200_greedy_argmax with a caller workspace", rows, vocab, need,
                (int)MVB200_GREEDY_ARGMAX_STATIC_BYTES);
  unsigned long long* scratch = nullptr;
  if (int rc = static_argmax_scratch(&scratch)) return rc;
  // Nothing to clear: pass one overwrites every key pass two reads (rows x chunks of them), so what an
  // earlier call left behind is never seen.
  cudaStream_t s = (cudaStream_t)stream;
  note_kernel("greedy_argmax_static");
  switch (dtype) {
    case MVB200_BF16:
      return launch_argmax<__nv_bfloat16>(logits, row_stride, rows, vocab, scratch, out_index, out_value, index_offset, s);
    case MVB200_F16:
      return launch_argmax<__half>(logits, row_stride, rows, vocab, scratch, out_index, out_value, index_offset, s);
    default:
      return launch_argmax<float>(logits, row_stride, rows, vocab, scratch, out_index, out_value, index_offset, s);
  }
}
