// qk_rope_store.cu — q/k RMSNorm + rotary embedding + KV-cache store in one pass.
//
// Replaces, on the caller side of the attention layer, the sequence in QwenAttention.forward
// (reference minivllm/models/qwen_base.py:175-186):
//     q, _ = q_norm(q); k, _ = k_norm(k)              layers/layernorm.py:81-93      (Qwen3 only)
//     q, k = rotary_emb(positions, q, k)              layers/rotary_embedding.py:166-193, :78-84
//     store_kv_cache(k, v, k_cache, v_cache, slots)   layers/attention_backend.py:196-266 (first thing
//                                                     Attention.forward does, attention.py:228-243)
// which is 2 x (pow, mean, add, rsqrt, mul, mul, cast) + 2 x (cast, chunk, 4 mul, add, sub, cat, cast)
// + the scatter: ~30 elementwise launches and as many round trips over q and k.  Here every
// (token, head) row is read once and written once; K goes to its cache slot straight from registers.
//
// HBM-bound.  Algorithmic bytes per token: (H + 2 KVH) D e read + (H + KVH) D e written back
// + 2 KVH D e into the cache (+ 8 B position, 4 B slot; the D/2 cos and sin floats of a token are
// shared by all of its heads and stay in L1/L2).
//
// Arithmetic follows the reference's fp32 operation order and rounding points so that the result
// matches its CPU execution bit for bit up to the summation order inside mean():
//   xf = float(x); inv = 1 / sqrt(sum(xf^2) / D + eps); xn = T((xf * inv) * float(w))
//   y1 = T(x1*cos - x2*sin), y2 = T(x2*cos + x1*sin) on float(xn) halves, products rounded
//   separately (no FMA contraction: __fmul_rn / __fadd_rn / __fsub_rn).
// One warp per row; lane l owns element pairs (2p, 2p+1) of both halves for p = l, l+32, ...
#include <stdlib.h>

#include "common.cuh"

namespace mvb200 {

namespace rope {

constexpr int kMaxIter = 2;  // head_dim <= 256

template <typename T> struct Pair;  // two adjacent elements
template <> struct Pair<float> {
  using V = float2;
  static MVB_DEV void unpack(V v, float& a, float& b) { a = v.x; b = v.y; }
  static MVB_DEV V pack(float a, float b) { return make_float2(a, b); }
};
template <> struct Pair<__half> {
  using V = __half2;
  static MVB_DEV void unpack(V v, float& a, float& b) { a = __low2float(v); b = __high2float(v); }
  static MVB_DEV V pack(float a, float b) { return __halves2half2(__float2half_rn(a), __float2half_rn(b)); }
};
template <> struct Pair<__nv_bfloat16> {
  using V = __nv_bfloat162;
  static MVB_DEV void unpack(V v, float& a, float& b) { a = __low2float(v); b = __high2float(v); }
  static MVB_DEV V pack(float a, float b) { return __halves2bfloat162(__float2bfloat16_rn(a), __float2bfloat16_rn(b)); }
};

struct RopeParams {
  const void* q; int64_t q_stride;
  const void* k; int64_t k_stride;
  const void* v; int64_t v_stride;
  void* q_out; int64_t qo_stride;
  void* k_out; int64_t ko_stride;  // may be null
  const int64_t* positions;
  const float* cos_cache;
  const float* sin_cache;
  int64_t max_pos;
  const void* q_weight;  // null: no norm
  const void* k_weight;
  float eps;
  const int32_t* slots;  // null: no store
  void* k_cache;
  void* v_cache;
  int64_t num_slots;
  int64_t total_rows;
  int32_t H, KVH, D;
};

template <typename T>
__global__ void __launch_bounds__(256)
qk_rope_store_kernel(const RopeParams p) {
  using P = Pair<T>;
  using V = typename P::V;
  pdl_wait();  // q/k/v come from the qkv GEMM; the cache may still be read by the previous step
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31;
  const int warps_per_cta = blockDim.x >> 5;
  const int64_t warp0 = (int64_t)blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
  const int64_t warp_stride = (int64_t)gridDim.x * warps_per_cta;
  const int rows_per_tok = p.H + 2 * p.KVH;
  const int half = p.D >> 1;
  const int pairs = half >> 1;  // element pairs per half

  for (int64_t row = warp0; row < p.total_rows; row += warp_stride) {
    const int64_t tok = row / rows_per_tok;
    const int j = (int)(row - tok * rows_per_tok);
    const bool is_q = j < p.H;
    const bool is_k = !is_q && j < p.H + p.KVH;
    const int head = is_q ? j : (is_k ? j - p.H : j - p.H - p.KVH);
    int64_t slot = -1;
    if (!is_q && p.slots) {
      slot = p.slots[tok];
      if (slot >= p.num_slots) slot = -1;
    }
    if (!is_q && !is_k) {
      // ---- V row: plain copy into the cache ----
      if (slot >= 0) {
        const V* src = reinterpret_cast<const V*>(static_cast<const T*>(p.v) + tok * p.v_stride + (int64_t)head * p.D);
        V* dst = reinterpret_cast<V*>(static_cast<T*>(p.v_cache) + (slot * p.KVH + head) * p.D);
        for (int i = lane; i < half; i += 32) dst[i] = src[i];
      }
      continue;
    }
    const T* src = is_q ? static_cast<const T*>(p.q) + tok * p.q_stride + (int64_t)head * p.D
                        : static_cast<const T*>(p.k) + tok * p.k_stride + (int64_t)head * p.D;
    const V* s1 = reinterpret_cast<const V*>(src);
    const V* s2 = reinterpret_cast<const V*>(src + half);
    float a[kMaxIter][2], b[kMaxIter][2];  // first-half / second-half values
    float ss = 0.f;
#pragma unroll
    for (int it = 0; it < kMaxIter; ++it) {
      const int pi = lane + 32 * it;
      a[it][0] = a[it][1] = b[it][0] = b[it][1] = 0.f;
      if (pi < pairs) {
        P::unpack(s1[pi], a[it][0], a[it][1]);
        P::unpack(s2[pi], b[it][0], b[it][1]);
      }
    }
    const T* w = static_cast<const T*>(is_q ? p.q_weight : p.k_weight);
    if (w) {
#pragma unroll
      for (int it = 0; it < kMaxIter; ++it)
        ss += __fmul_rn(a[it][0], a[it][0]) + __fmul_rn(a[it][1], a[it][1]) +
              __fmul_rn(b[it][0], b[it][0]) + __fmul_rn(b[it][1], b[it][1]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
      const float var = __fdiv_rn(ss, (float)p.D);
      const float inv = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, p.eps)));
      const V* w1 = reinterpret_cast<const V*>(w);
      const V* w2 = reinterpret_cast<const V*>(w + half);
#pragma unroll
      for (int it = 0; it < kMaxIter; ++it) {
        const int pi = lane + 32 * it;
        if (pi < pairs) {
          float wa0, wa1, wb0, wb1;
          P::unpack(w1[pi], wa0, wa1);
          P::unpack(w2[pi], wb0, wb1);
          // round to the model dtype like RMSNorm's .to(orig_dtype), then back to fp32 for the rotation
          P::unpack(P::pack(__fmul_rn(__fmul_rn(a[it][0], inv), wa0), __fmul_rn(__fmul_rn(a[it][1], inv), wa1)),
                    a[it][0], a[it][1]);
          P::unpack(P::pack(__fmul_rn(__fmul_rn(b[it][0], inv), wb0), __fmul_rn(__fmul_rn(b[it][1], inv), wb1)),
                    b[it][0], b[it][1]);
        }
      }
    }
    int64_t pos = p.positions[tok];
    pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);  // the reference would raise IndexError
    const float2* cs = reinterpret_cast<const float2*>(p.cos_cache + pos * half);
    const float2* sn = reinterpret_cast<const float2*>(p.sin_cache + pos * half);
    T* dst = is_q ? static_cast<T*>(p.q_out) + tok * p.qo_stride + (int64_t)head * p.D
                  : (p.k_out ? static_cast<T*>(p.k_out) + tok * p.ko_stride + (int64_t)head * p.D : nullptr);
    T* cdst = (is_k && slot >= 0) ? static_cast<T*>(p.k_cache) + (slot * p.KVH + head) * p.D : nullptr;
#pragma unroll
    for (int it = 0; it < kMaxIter; ++it) {
      const int pi = lane + 32 * it;
      if (pi < pairs) {
        const float2 c = __ldg(cs + pi), s = __ldg(sn + pi);
        const V y1 = P::pack(__fsub_rn(__fmul_rn(a[it][0], c.x), __fmul_rn(b[it][0], s.x)),
                             __fsub_rn(__fmul_rn(a[it][1], c.y), __fmul_rn(b[it][1], s.y)));
        const V y2 = P::pack(__fadd_rn(__fmul_rn(b[it][0], c.x), __fmul_rn(a[it][0], s.x)),
                             __fadd_rn(__fmul_rn(b[it][1], c.y), __fmul_rn(a[it][1], s.y)));
        if (dst) {
          reinterpret_cast<V*>(dst)[pi] = y1;
          reinterpret_cast<V*>(dst + half)[pi] = y2;
        }
        if (cdst) {
          reinterpret_cast<V*>(cdst)[pi] = y1;
          reinterpret_cast<V*>(cdst + half)[pi] = y2;
        }
      }
    }
  }
}

// ---- vectorised path: 16-bit dtypes, head_dim 128, 16-byte aligned rows ---------------------------
// The pair kernel above keeps one 256-byte row per warp in flight with 4-byte accesses; at prefill sizes
// (cfg3: 8192 tokens x 48 rows) that is latency-bound at ~1.6 TB/s.  Here 8 lanes own a row (16 bytes of
// each half per lane) and a warp works on 4 rows at a time: 1 KiB in flight per warp, 128-bit accesses.
// Same arithmetic and rounding points; only the order of the 128-term sum inside mean() differs.
template <typename T> struct Vec8;
template <> struct Vec8<__nv_bfloat16> {
  static MVB_DEV void unpack(const uint4& r, float* f) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { f[2 * i] = __uint_as_float(w[i] << 16); f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u); }
  }
  static MVB_DEV uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { __nv_bfloat162 v = __halves2bfloat162(__float2bfloat16_rn(f[2 * i]), __float2bfloat16_rn(f[2 * i + 1])); w[i] = *reinterpret_cast<uint32_t*>(&v); }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <> struct Vec8<__half> {
  static MVB_DEV void unpack(const uint4& r, float* f) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&w[i])); f[2 * i] = v.x; f[2 * i + 1] = v.y; }
  }
  static MVB_DEV uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { __half2 v = __halves2half2(__float2half_rn(f[2 * i]), __float2half_rn(f[2 * i + 1])); w[i] = *reinterpret_cast<uint32_t*>(&v); }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

template <typename T>
__global__ void __launch_bounds__(256)
qk_rope_store_vec_kernel(const RopeParams p) {
  constexpr int D = 128, HALF = 64;
  pdl_wait();
  pdl_launch_dependents();
  const int lane = threadIdx.x & 31;
  const int li = lane & 7;       // 16-byte chunk of each half owned by this lane
  const int rsub = lane >> 3;    // which of the warp's 4 rows
  const int warps_per_cta = blockDim.x >> 5;
  const int64_t warp0 = (int64_t)blockIdx.x * warps_per_cta + (threadIdx.x >> 5);
  const int64_t warp_stride = (int64_t)gridDim.x * warps_per_cta;
  const int rows_per_tok = p.H + 2 * p.KVH;
  const int64_t groups = (p.total_rows + 3) >> 2;

  for (int64_t grp = warp0; grp < groups; grp += warp_stride) {
    const int64_t row = grp * 4 + rsub;
    const bool valid = row < p.total_rows;
    const int64_t tok = valid ? row / rows_per_tok : 0;
    const int j = valid ? (int)(row - tok * rows_per_tok) : 0;
    const bool is_q = j < p.H;
    const bool is_k = !is_q && j < p.H + p.KVH;
    const int head = is_q ? j : (is_k ? j - p.H : j - p.H - p.KVH);
    int64_t slot = -1;
    if (valid && !is_q && p.slots) {
      slot = p.slots[tok];
      if (slot >= p.num_slots) slot = -1;
    }
    const bool is_v = !is_q && !is_k;
    const T* src = is_q ? static_cast<const T*>(p.q) + tok * p.q_stride + (int64_t)head * D
                        : (is_k ? static_cast<const T*>(p.k) + tok * p.k_stride + (int64_t)head * D
                                : static_cast<const T*>(p.v) + tok * p.v_stride + (int64_t)head * D);
    uint4 ra = make_uint4(0, 0, 0, 0), rb = make_uint4(0, 0, 0, 0);
    if (valid && !(is_v && slot < 0)) {
      ra = *reinterpret_cast<const uint4*>(src + li * 8);
      rb = *reinterpret_cast<const uint4*>(src + HALF + li * 8);
    }
    if (is_v) {  // plain copy into the cache (no shuffles below depend on these lanes' data)
      if (valid && slot >= 0) {
        T* dst = static_cast<T*>(p.v_cache) + (slot * p.KVH + head) * D;
        *reinterpret_cast<uint4*>(dst + li * 8) = ra;
        *reinterpret_cast<uint4*>(dst + HALF + li * 8) = rb;
      }
    }
    float a[8], b[8];
    Vec8<T>::unpack(ra, a);
    Vec8<T>::unpack(rb, b);
    const T* w = static_cast<const T*>(is_q ? p.q_weight : p.k_weight);
    if (p.q_weight) {  // uniform: q and k weights come together; every lane takes part in the shuffles
      float ss = 0.f;
#pragma unroll
      for (int e = 0; e < 8; ++e) ss += __fmul_rn(a[e], a[e]) + __fmul_rn(b[e], b[e]);
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
      if (valid && !is_v) {
        const float var = __fdiv_rn(ss, (float)D);
        const float inv = __fdiv_rn(1.f, __fsqrt_rn(__fadd_rn(var, p.eps)));
        float wa[8], wb[8];
        Vec8<T>::unpack(*reinterpret_cast<const uint4*>(w + li * 8), wa);
        Vec8<T>::unpack(*reinterpret_cast<const uint4*>(w + HALF + li * 8), wb);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          a[e] = __fmul_rn(__fmul_rn(a[e], inv), wa[e]);
          b[e] = __fmul_rn(__fmul_rn(b[e], inv), wb[e]);
        }
        // RMSNorm's .to(orig_dtype), then back to fp32 for the rotation
        Vec8<T>::unpack(Vec8<T>::pack(a), a);
        Vec8<T>::unpack(Vec8<T>::pack(b), b);
      }
    }
    if (!valid || is_v) continue;
    int64_t pos = p.positions[tok];
    pos = pos < 0 ? 0 : (pos >= p.max_pos ? p.max_pos - 1 : pos);
    const float4* cs = reinterpret_cast<const float4*>(p.cos_cache + pos * HALF + li * 8);
    const float4* sn = reinterpret_cast<const float4*>(p.sin_cache + pos * HALF + li * 8);
    const float4 c0 = __ldg(cs), c1 = __ldg(cs + 1), s0 = __ldg(sn), s1 = __ldg(sn + 1);
    const float c[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
    const float sv[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
    float y1[8], y2[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      y1[e] = __fsub_rn(__fmul_rn(a[e], c[e]), __fmul_rn(b[e], sv[e]));
      y2[e] = __fadd_rn(__fmul_rn(b[e], c[e]), __fmul_rn(a[e], sv[e]));
    }
    const uint4 o1 = Vec8<T>::pack(y1), o2 = Vec8<T>::pack(y2);
    T* dst = is_q ? static_cast<T*>(p.q_out) + tok * p.qo_stride + (int64_t)head * D
                  : (p.k_out ? static_cast<T*>(p.k_out) + tok * p.ko_stride + (int64_t)head * D : nullptr);
    if (dst) {
      *reinterpret_cast<uint4*>(dst + li * 8) = o1;
      *reinterpret_cast<uint4*>(dst + HALF + li * 8) = o2;
    }
    if (is_k && slot >= 0) {
      T* cdst = static_cast<T*>(p.k_cache) + (slot * p.KVH + head) * D;
      *reinterpret_cast<uint4*>(cdst + li * 8) = o1;
      *reinterpret_cast<uint4*>(cdst + HALF + li * 8) = o2;
    }
  }
}

template <typename T>
int launch_rope_vec(const RopeParams& p, cudaStream_t stream) {
  const int threads = 256, wpc = threads / 32;
  const int64_t groups = (p.total_rows + 3) / 4;
  int64_t blocks = (groups + wpc - 1) / wpc;
  const int64_t cap = 148LL * 16;
  if (blocks > cap) blocks = cap;
  cudaError_t e = launch_pdl(qk_rope_store_vec_kernel<T>, dim3((unsigned)blocks), dim3(threads), 0, stream, p);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "qk_norm_rope_store: %s", cudaGetErrorString(e));
  return check_launch("qk_norm_rope_store");
}

template <typename T>
int launch_rope(const RopeParams& p, cudaStream_t stream) {
  const int threads = 256, wpc = threads / 32;
  int64_t blocks = (p.total_rows + wpc - 1) / wpc;
  const int64_t cap = 148LL * 16;  // 8 CTAs of 256 threads fit an SM; two waves, then grid-stride
  if (blocks > cap) blocks = cap;
  cudaError_t e = launch_pdl(qk_rope_store_kernel<T>, dim3((unsigned)blocks), dim3(threads), 0, stream, p);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "qk_norm_rope_store: %s", cudaGetErrorString(e));
  return check_launch("qk_norm_rope_store");
}

}  // namespace rope
}  // namespace mvb200

extern "C" int mvb200_qk_norm_rope_store(
    const void* q, int64_t q_row_stride, const void* k, int64_t k_row_stride, const void* v,
    int64_t v_row_stride, void* q_out, int64_t q_out_row_stride, void* k_out,
    int64_t k_out_row_stride, const int64_t* positions, const float* cos_cache,
    const float* sin_cache, int64_t max_positions, const void* q_norm_weight,
    const void* k_norm_weight, float eps, const int32_t* slot_mapping, void* k_cache, void* v_cache,
    int64_t num_slots, int32_t num_tokens, int32_t num_heads, int32_t num_kv_heads,
    int32_t head_dim, int32_t dtype, void* stream) {
  using namespace mvb200;
  using namespace mvb200::rope;
  if (num_tokens < 0 || num_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0 || max_positions <= 0)
    return fail(MVB200_EINVAL, "qk_norm_rope_store: bad sizes");
  if (head_dim % 4 || head_dim > 128 * kMaxIter)
    return fail(MVB200_EINVAL, "qk_norm_rope_store: head_dim %d must be a multiple of 4 and <= %d",
                head_dim, 128 * kMaxIter);
  if (num_tokens == 0) return MVB200_OK;
  if (!q || !k || !v || !q_out || !positions || !cos_cache || !sin_cache)
    return fail(MVB200_EINVAL, "qk_norm_rope_store: null pointer");
  if ((q_norm_weight == nullptr) != (k_norm_weight == nullptr))
    return fail(MVB200_EINVAL, "qk_norm_rope_store: q and k norm weights must be given together");
  if (slot_mapping && (!k_cache || !v_cache || num_slots <= 0))
    return fail(MVB200_EINVAL, "qk_norm_rope_store: slot_mapping given without a cache");
  const int eb = dtype == MVB200_F32 ? 4 : ((dtype == MVB200_F16 || dtype == MVB200_BF16) ? 2 : 0);
  if (!eb) return fail(MVB200_EINVAL, "qk_norm_rope_store: unknown dtype %d", dtype);
  // pair-vector accesses: every base pointer and row stride must be aligned to two elements
  const int64_t strides[] = {q_row_stride, k_row_stride, v_row_stride, q_out_row_stride,
                             k_out ? k_out_row_stride : 0};
  for (int64_t s : strides)
    if (s % 2) return fail(MVB200_EINVAL, "qk_norm_rope_store: row strides must be even");
  const void* ptrs[] = {q, k, v, q_out, k_out, q_norm_weight, k_norm_weight, k_cache, v_cache};
  for (const void* ptr : ptrs)
    if (ptr && (reinterpret_cast<uintptr_t>(ptr) % (2 * eb)))
      return fail(MVB200_EINVAL, "qk_norm_rope_store: pointers must be aligned to %d bytes", 2 * eb);
  if ((reinterpret_cast<uintptr_t>(cos_cache) | reinterpret_cast<uintptr_t>(sin_cache)) % 8)
    return fail(MVB200_EINVAL, "qk_norm_rope_store: cos/sin tables must be 8-byte aligned");

  RopeParams p{q, q_row_stride, k, k_row_stride, v, v_row_stride, q_out, q_out_row_stride, k_out,
               k_out_row_stride, positions, cos_cache, sin_cache, max_positions, q_norm_weight,
               k_norm_weight, eps, slot_mapping, k_cache, v_cache, num_slots,
               (int64_t)num_tokens * (num_heads + 2 * num_kv_heads), num_heads, num_kv_heads, head_dim};
  note_kernel("qk_norm_rope_store");
  cudaStream_t s = (cudaStream_t)stream;
  // 128-bit path: 16-bit elements, head_dim 128, every row start 16-byte aligned
  bool vec_ok = eb == 2 && head_dim == 128 && tuning_env("MVB200_ROPE_PAIR") == nullptr;
  for (int64_t st : strides) vec_ok = vec_ok && (st % 8 == 0);
  for (const void* ptr : ptrs) vec_ok = vec_ok && (reinterpret_cast<uintptr_t>(ptr) % 16 == 0);
  vec_ok = vec_ok && ((reinterpret_cast<uintptr_t>(cos_cache) | reinterpret_cast<uintptr_t>(sin_cache)) % 16 == 0);
  if (vec_ok) return dtype == MVB200_F16 ? launch_rope_vec<__half>(p, s) : launch_rope_vec<__nv_bfloat16>(p, s);
  if (dtype == MVB200_F32) return launch_rope<float>(p, s);
  if (dtype == MVB200_F16) return launch_rope<__half>(p, s);
  return launch_rope<__nv_bfloat16>(p, s);
}
