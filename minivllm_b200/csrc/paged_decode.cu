// paged_decode.cu — split-KV paged GQA decode attention for sm_100a.
//
// Replaces flash_attn_with_kvcache at reference minivllm/models/layers/attention.py:439-447 and
// the eager PageAttention.__call__ (minivllm/models/layers/page_attention.py:35-133).
//
// Fast path (16-bit K/V, D in {64,128}, block_size % 64 == 0, group <= 16):
//   grid = (KVH, num_splits, B); one CTA streams the 64-token tiles of ONE kv head of ONE
//   sequence split.  A producer warp turns block-table entries into TMA coordinates and issues
//   cp.async.bulk.tensor loads of whole K and V tiles (64 tokens x D, 128B-swizzled) into a
//   multi-stage shared-memory ring guarded by full/empty mbarriers.  Four consumer warps each
//   own 16 tokens of a tile: Q(group padded to 16 rows) x K^T and P x V run on mma.sync
//   m16n8k16 with ldmatrix-fed fragments (K/V are read from HBM once per kv head, not once per
//   query head), online softmax in fp32 with quad shuffles, exp2 with the scale folded in.
//   Warps are merged through shared memory at the end; with num_splits > 1 each CTA writes a
//   normalised partial + log-sum-exp and decode_reduce_kernel combines them.
//   HBM-bound: algorithmic bytes = ctx * D * 2 (K,V) * elem per (sequence, kv head).
//
// Generic path (fp32 caches, any D <= 256, any block size, any group): one CTA per (b, h),
//   CUDA-core dot products.  Used for the reference's CPU-sized unit-test shapes.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace mvb200 {

// =============================================================================================
// fast path
// =============================================================================================

constexpr int kTileTokens = 64;
constexpr int kConsumerWarps = 4;
constexpr int kDecodeThreads = (kConsumerWarps + 1) * 32;

struct DecodeParams {
  const void* q;
  int64_t q_row_stride;  // elements between tokens
  const int32_t* block_tables;
  int64_t bt_row_stride;
  const int32_t* context_lens;
  void* out;
  int64_t out_row_stride;
  float* part_o;    // [B, H, S, D]
  float* part_lse;  // [B, H, S]  (log2 domain)
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t group;
  int32_t max_blocks;      // block-table width (MB)
  int32_t tiles_per_block;  // block_size / 64
  int32_t num_blocks;       // NB (also the out-of-bounds TMA coordinate -> zero fill)
  int32_t num_splits;
  int32_t stages;
  float scale_log2;
};

template <typename T, int D>
__global__ void __launch_bounds__(kDecodeThreads)
paged_decode_kernel(const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const DecodeParams p) {
  constexpr int kHalves = D / 64;                       // 128-byte column groups per row
  constexpr int kHalfBytes = kTileTokens * 128;         // one [64 tokens][128 B] swizzled slab
  constexpr int kTileBytes = kHalves * kHalfBytes;      // K (or V) tile
  constexpr int kStageBytes = 2 * kTileBytes;           // K + V
  constexpr int kKSteps = D / 16;

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem base is only guaranteed 16B aligned: round up to the swizzle atom
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int stages = p.stages;
  const uint32_t bar_base = smem_base + stages * kStageBytes;  // full[stages], empty[stages]

  const int kvh = blockIdx.x;
  const int split = blockIdx.y;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int G = p.group;

  // prologue that needs nothing from earlier kernels runs before the dependency wait (PDL)
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(bar_base + 8 * s, 1);                           // full: producer's expect_tx arrive
      mbar_init(bar_base + 8 * (stages + s), kConsumerWarps);   // empty: one arrive per warp
    }
    fence_mbar_init();
  }
  if (threadIdx.x == kConsumerWarps * 32) {
    tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v);
  }
  __syncthreads();
  pdl_wait();               // q, context_lens, block tables and the freshly stored K/V are visible from here
  pdl_launch_dependents();  // the split reduction (or the next layer's kernel) may start its own prologue

  const int ctx = p.context_lens[b];
  const int n_tiles = ctx > 0 ? (ctx + kTileTokens - 1) / kTileTokens : 0;
  const int tps = (n_tiles + p.num_splits - 1) / p.num_splits;
  const int t0 = split * tps;
  const int t1 = min(n_tiles, t0 + tps);

  T* out = reinterpret_cast<T*>(p.out);
  if (t1 <= t0) {
    // empty split (or padded row with context_len == 0): publish "no contribution"
    if (p.num_splits > 1) {
      for (int r = threadIdx.x; r < G; r += blockDim.x) {
        const int h = kvh * G + r;
        p.part_lse[((int64_t)b * p.num_heads + h) * p.num_splits + split] = -INFINITY;
      }
    } else {
      for (int i = threadIdx.x; i < G * D; i += blockDim.x)
        out[(int64_t)b * p.out_row_stride + (int64_t)kvh * G * D + i] = from_f32<T>(0.f);
    }
    return;
  }

  // per-warp running state (consumer warps only)
  float o_acc[D / 8][4];
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};

  if (warp == kConsumerWarps) {
    // ------------------------------ producer warp ------------------------------
    const int n = t1 - t0;
    const int32_t* bt = p.block_tables + (int64_t)b * p.bt_row_stride;
    for (int base = 0; base < n; base += 32) {
      // 32 block-table entries per coalesced load, then one lane issues the copies
      int my_bid = p.num_blocks;  // out-of-bounds coordinate: TMA fills the tile with zeros
      int my_sub = 0;
      if (base + lane < n) {
        const int t = t0 + base + lane;
        const int j = min(t / p.tiles_per_block, p.max_blocks - 1);  // page_attention.py:83-84
        const int v = __ldg(bt + j);
        if (v >= 0 && v < p.num_blocks) my_bid = v;
        my_sub = (t % p.tiles_per_block) * kTileTokens;
      }
      const int cnt = min(32, n - base);
      for (int i = 0; i < cnt; ++i) {
        const int bid = __shfl_sync(0xffffffffu, my_bid, i);
        const int sub = __shfl_sync(0xffffffffu, my_sub, i);
        if (lane == 0) {
          const int it = base + i;
          const int s = it % stages;
          const uint32_t full = bar_base + 8 * s;
          const uint32_t empty = bar_base + 8 * (stages + s);
          mbar_wait(empty, ((it / stages) & 1) ^ 1);
          mbar_arrive_expect_tx(full, kStageBytes);
          const uint32_t kdst = smem_base + s * kStageBytes;
          const uint32_t vdst = kdst + kTileBytes;
#pragma unroll
          for (int h = 0; h < kHalves; ++h) {
            tma_load_4d(kdst + h * kHalfBytes, &tmap_k, full, h * 64, kvh, sub, bid);
            tma_load_4d(vdst + h * kHalfBytes, &tmap_v, full, h * 64, kvh, sub, bid);
          }
        }
        __syncwarp();
      }
    }
  } else {
    // ------------------------------ consumer warps -----------------------------
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;

    // Q fragments (A operand, 16 x D, rows >= group are zero), straight from global
    uint32_t qf[kKSteps][4];
    {
      const T* qb = reinterpret_cast<const T*>(p.q) + (int64_t)b * p.q_row_stride + (int64_t)kvh * G * D;
      const int g = lane >> 2, t = lane & 3;
#pragma unroll
      for (int kk = 0; kk < kKSteps; ++kk) {
        const int c = kk * 16 + 2 * t;
        qf[kk][0] = g < G ? *reinterpret_cast<const uint32_t*>(qb + g * D + c) : 0u;
        qf[kk][1] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(qb + (g + 8) * D + c) : 0u;
        qf[kk][2] = g < G ? *reinterpret_cast<const uint32_t*>(qb + g * D + c + 8) : 0u;
        qf[kk][3] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(qb + (g + 8) * D + c + 8) : 0u;
      }
    }

    const int n = t1 - t0;
    // ldmatrix lane geometry (see DESIGN.md "decode fragment map")
    const int krow = 16 * warp + (lane & 7) + ((lane >> 4) << 3);
    const int ksel = (lane >> 3) & 1;
    const int vrow = 16 * warp + (lane & 7) + (((lane >> 3) & 1) << 3);
    const int vsel = lane >> 4;

    for (int it = 0; it < n; ++it) {
      const int s = it % stages;
      const uint32_t full = bar_base + 8 * s;
      const uint32_t empty = bar_base + 8 * (stages + s);
      const uint32_t kbase = smem_base + s * kStageBytes;
      const uint32_t vbase = kbase + kTileBytes;
      const int tile_tok0 = (t0 + it) * kTileTokens;
      const bool partial = tile_tok0 + kTileTokens > ctx;

      mbar_wait(full, (it / stages) & 1);

      // ---- S = Q K^T for this warp's 16 tokens ----
      float sacc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int kk = 0; kk < kKSteps; ++kk) {
        const int cidx = kk * 2 + ksel;
        const uint32_t addr = kbase + (cidx >> 3) * kHalfBytes + krow * 128 +
                              (((cidx & 7) ^ (krow & 7)) << 4);
        uint32_t kf[4];
        ldmatrix_x4(kf, addr);
        Mma16816<T>::run(sacc[0], qf[kk], kf[0], kf[1]);
        Mma16816<T>::run(sacc[1], qf[kk], kf[2], kf[3]);
      }

      if (partial) {
        // tokens >= ctx: score -> -inf (select, K garbage may be NaN) and V rows -> 0 (0*NaN)
        const int valid = ctx - tile_tok0;  // 1..63 valid tokens in this tile
        const int c = 16 * warp + 2 * (lane & 3);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          if (c + 8 * j >= valid) sacc[j][0] = sacc[j][2] = -INFINITY;
          if (c + 8 * j + 1 >= valid) sacc[j][1] = sacc[j][3] = -INFINITY;
        }
        const int first_bad = max(valid, 16 * warp);
        for (int r = first_bad; r < 16 * warp + 16; ++r) {
          if (lane < kHalves * 8) {
            uint8_t* dst = smem_gen + (vbase - smem_base) + (lane >> 3) * kHalfBytes + r * 128 + (lane & 7) * 16;
            *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
          }
        }
        fence_proxy_async_smem();  // generic writes before the next TMA refill of this slab
        __syncwarp();
      }

      // ---- online softmax (log2 domain) ----
      float alpha[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float mx = fmaxf(fmaxf(sacc[0][2 * r], sacc[0][2 * r + 1]),
                         fmaxf(sacc[1][2 * r], sacc[1][2 * r + 1]));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float m_new = fmaxf(m_run[r], mx * p.scale_log2);
        const float m_use = m_new == -INFINITY ? 0.f : m_new;
        alpha[r] = exp2f(m_run[r] - m_use);
        m_run[r] = m_new;
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          sacc[j][2 * r] = exp2f(sacc[j][2 * r] * p.scale_log2 - m_use);
          sacc[j][2 * r + 1] = exp2f(sacc[j][2 * r + 1] * p.scale_log2 - m_use);
          sum += sacc[j][2 * r] + sacc[j][2 * r + 1];
        }
        l_run[r] = l_run[r] * alpha[r] + sum;
      }
#pragma unroll
      for (int i = 0; i < D / 8; ++i) {
        o_acc[i][0] *= alpha[0];
        o_acc[i][1] *= alpha[0];
        o_acc[i][2] *= alpha[1];
        o_acc[i][3] *= alpha[1];
      }

      // ---- O += P V ----
      uint32_t pf[4];
      pf[0] = Mma16816<T>::pack(sacc[0][0], sacc[0][1]);
      pf[1] = Mma16816<T>::pack(sacc[0][2], sacc[0][3]);
      pf[2] = Mma16816<T>::pack(sacc[1][0], sacc[1][1]);
      pf[3] = Mma16816<T>::pack(sacc[1][2], sacc[1][3]);
#pragma unroll
      for (int np = 0; np < D / 16; ++np) {
        const int cidx = np * 2 + vsel;
        const uint32_t addr = vbase + (cidx >> 3) * kHalfBytes + vrow * 128 +
                              (((cidx & 7) ^ (vrow & 7)) << 4);
        uint32_t vf[4];
        ldmatrix_x4_trans(vf, addr);
        Mma16816<T>::run(o_acc[2 * np], pf, vf[0], vf[1]);
        Mma16816<T>::run(o_acc[2 * np + 1], pf, vf[2], vf[3]);
      }

      __syncwarp();
      if (lane == 0) mbar_arrive(empty);
    }

    // finish row sums across the quad
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }
  }

  // ------------------------------ merge the four warps ------------------------------
  // every issued tile has been consumed, so the ring can be reused as scratch
  __syncthreads();
  constexpr int kRowF = D + 4;  // padded row (floats)
  float* mo = reinterpret_cast<float*>(smem_gen);                       // [4][16][kRowF]
  float* mm = mo + kConsumerWarps * 16 * kRowF;                         // [4][16]
  float* ml = mm + kConsumerWarps * 16;                                 // [4][16]
  if (warp < kConsumerWarps) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int row = g + 8 * r;
      if (row < G) {
        float* dst = mo + (warp * 16 + row) * kRowF;
#pragma unroll
        for (int i = 0; i < D / 8; ++i)
          *reinterpret_cast<float2*>(dst + i * 8 + 2 * t) = make_float2(o_acc[i][2 * r], o_acc[i][2 * r + 1]);
        if (t == 0) {
          mm[warp * 16 + row] = m_run[r];
          ml[warp * 16 + row] = l_run[r];
        }
      }
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < G * D; idx += blockDim.x) {
    const int row = idx / D, d = idx - row * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, mm[w * 16 + row]);
    const float Ms = M == -INFINITY ? 0.f : M;
    float L = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < kConsumerWarps; ++w) {
      const float sc = exp2f(mm[w * 16 + row] - Ms);
      L += ml[w * 16 + row] * sc;
      o += mo[(w * 16 + row) * kRowF + d] * sc;
    }
    const float val = L > 0.f ? o / L : 0.f;
    const int h = kvh * G + row;
    if (p.num_splits == 1) {
      out[(int64_t)b * p.out_row_stride + (int64_t)h * D + d] = from_f32<T>(val);
    } else {
      const int64_t ph = ((int64_t)b * p.num_heads + h) * p.num_splits + split;
      p.part_o[ph * D + d] = val;
      if (d == 0) p.part_lse[ph] = L > 0.f ? M + log2f(L) : -INFINITY;
    }
  }
}

// Combine the per-split partials: out = sum_s w_s o_s / sum_s w_s, w_s = 2^(lse_s - max lse).
// One CTA per (head, sequence); the four warps take interleaved splits so that the partial rows of
// different splits are in flight together (a serial walk over 128 splits costs more than the decode
// kernel itself at batch 1), lanes cover head_dim with coalesced 128-byte rows.
template <typename T>
__global__ void __launch_bounds__(128)
decode_reduce_kernel(const float* __restrict__ part_o, const float* __restrict__ part_lse,
                     T* __restrict__ out, int64_t out_row_stride, int num_heads, int head_dim,
                     int num_splits) {
  constexpr int kMaxPerLane = 8;  // head_dim <= 256
  __shared__ float sm_acc[4][256];
  __shared__ float sm_w[4], sm_max[4];
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t ph = ((int64_t)b * num_heads + h) * num_splits;
  pdl_wait();  // partials come from the decode kernel launched just before
  pdl_launch_dependents();
  float m = -INFINITY;
  for (int s = threadIdx.x; s < num_splits; s += 128) m = fmaxf(m, part_lse[ph + s]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (lane == 0) sm_max[warp] = m;
  __syncthreads();
  const float M = fmaxf(fmaxf(sm_max[0], sm_max[1]), fmaxf(sm_max[2], sm_max[3]));
  float acc[kMaxPerLane];
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) acc[i] = 0.f;
  float wsum = 0.f;
  if (M != -INFINITY) {
    for (int s = warp; s < num_splits; s += 4) {
      const float lse = part_lse[ph + s];
      const float w = lse == -INFINITY ? 0.f : exp2f(lse - M);  // empty split: its row is never read as data
      wsum += w;
      const float* row = part_o + (ph + s) * head_dim;
#pragma unroll
      for (int i = 0; i < kMaxPerLane; ++i) {
        const int d = lane + 32 * i;
        if (d < head_dim && w != 0.f) acc[i] += w * row[d];
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) sm_acc[warp][lane + 32 * i] = acc[i];
  if (lane == 0) sm_w[warp] = wsum;
  __syncthreads();
  const float W = sm_w[0] + sm_w[1] + sm_w[2] + sm_w[3];
  for (int d = threadIdx.x; d < head_dim; d += 128) {
    const float o = sm_acc[0][d] + sm_acc[1][d] + sm_acc[2][d] + sm_acc[3][d];
    out[(int64_t)b * out_row_stride + (int64_t)h * head_dim + d] = from_f32<T>(W > 0.f ? o / W : 0.f);
  }
}

// =============================================================================================
// generic path: one CTA (4 warps) per (sequence, query head); any dtype / head_dim / block size
// =============================================================================================
template <typename T>
__global__ void __launch_bounds__(128)
paged_decode_generic_kernel(const T* __restrict__ q, int64_t q_row_stride,
                            const T* __restrict__ k_cache, const T* __restrict__ v_cache,
                            const int32_t* __restrict__ block_tables, int64_t bt_row_stride,
                            const int32_t* __restrict__ context_lens, T* __restrict__ out,
                            int64_t out_row_stride, int num_heads, int num_kv_heads, int head_dim,
                            int block_size, int64_t num_blocks, int max_blocks, float scale_log2) {
  constexpr int kMaxPerLane = 8;  // head_dim <= 256
  const int h = blockIdx.x, b = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kvh = h / (num_heads / num_kv_heads);
  const int ctx = context_lens[b];
  __shared__ float sm_o[4][256];
  __shared__ float sm_m[4], sm_l[4];

  float qr[kMaxPerLane], acc[kMaxPerLane];
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) {
    const int d = lane + 32 * i;
    qr[i] = d < head_dim ? to_f32(q[(int64_t)b * q_row_stride + (int64_t)h * head_dim + d]) : 0.f;
    acc[i] = 0.f;
  }
  float m = -INFINITY, l = 0.f;
  const int64_t tok_stride = (int64_t)num_kv_heads * head_dim;
  for (int pos = warp; pos < ctx; pos += 4) {
    const int j = min(pos / block_size, max_blocks - 1);
    const int bid = block_tables[(int64_t)b * bt_row_stride + j];
    const bool ok = bid >= 0 && bid < num_blocks;  // invalid page: zero key/value row, score 0
    const int64_t base = ((int64_t)(ok ? bid : 0) * block_size + pos % block_size) * tok_stride +
                         (int64_t)kvh * head_dim;
    float kr[kMaxPerLane], dot = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i) {
      const int d = lane + 32 * i;
      kr[i] = (ok && d < head_dim) ? to_f32(k_cache[base + d]) : 0.f;
      dot += qr[i] * kr[i];
    }
    dot = warp_sum(dot) * scale_log2;
    const float m_new = fmaxf(m, dot);
    const float a = exp2f(m - m_new);
    const float pw = exp2f(dot - m_new);
    l = l * a + pw;
#pragma unroll
    for (int i = 0; i < kMaxPerLane; ++i) {
      const int d = lane + 32 * i;
      const float vv = (ok && d < head_dim) ? to_f32(v_cache[base + d]) : 0.f;
      acc[i] = acc[i] * a + pw * vv;
    }
    m = m_new;
  }
#pragma unroll
  for (int i = 0; i < kMaxPerLane; ++i) sm_o[warp][lane + 32 * i] = acc[i];
  if (lane == 0) {
    sm_m[warp] = m;
    sm_l[warp] = l;
  }
  __syncthreads();
  for (int d = threadIdx.x; d < head_dim; d += blockDim.x) {
    float M = fmaxf(fmaxf(sm_m[0], sm_m[1]), fmaxf(sm_m[2], sm_m[3]));
    const float Ms = M == -INFINITY ? 0.f : M;
    float L = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float sc = exp2f(sm_m[w] - Ms);
      L += sm_l[w] * sc;
      o += sm_o[w][d] * sc;
    }
    out[(int64_t)b * out_row_stride + (int64_t)h * head_dim + d] = from_f32<T>(L > 0.f ? o / L : 0.f);
  }
}

// =============================================================================================
// host launchers
// =============================================================================================

template <typename T, int D>
static int launch_fast(const DecodeParams& p, const void* k_cache, const void* v_cache,
                       int block_size, int batch, cudaStream_t stream) {
  CUtensorMap mk, mv;
  const uint64_t dims[4] = {(uint64_t)D, (uint64_t)p.num_kv_heads, (uint64_t)block_size,
                            (uint64_t)p.num_blocks};
  const uint64_t strides[4] = {0, (uint64_t)D * 2, (uint64_t)p.num_kv_heads * D * 2,
                               (uint64_t)block_size * p.num_kv_heads * D * 2};
  const uint32_t box[4] = {64, 1, (uint32_t)kTileTokens, 1};
  int rc = make_tensor_map(&mk, k_cache, 2, 4, dims, strides, box, true);
  if (rc) return rc;
  rc = make_tensor_map(&mv, v_cache, 2, 4, dims, strides, box, true);
  if (rc) return rc;

  constexpr int kStageBytes = 2 * (D / 64) * kTileTokens * 128;
  const int merge_bytes = (kConsumerWarps * 16 * (D + 4) + 2 * kConsumerWarps * 16) * 4;
  int smem = p.stages * kStageBytes + 2 * p.stages * 8;
  if (smem < merge_bytes) smem = merge_bytes;
  smem += 1024;  // alignment slack
  auto kern = paged_decode_kernel<T, D>;
  static SmemOptIn optin;
  rc = optin_dynamic_smem((const void*)kern, smem, &optin, "paged_decode");
  if (rc) return rc;
  dim3 grid(p.num_kv_heads, p.num_splits, batch);
  cudaError_t le = launch_pdl(kern, grid, dim3(kDecodeThreads), (size_t)smem, stream, mk, mv, p);
  if (le != cudaSuccess)
    return fail(MVB200_ECUDA, "paged_decode_kernel: %s", cudaGetErrorString(le));
  rc = check_launch("paged_decode_kernel");
  if (rc) return rc;
  if (p.num_splits > 1) {
    dim3 rgrid(p.num_heads, batch);
    le = launch_pdl(decode_reduce_kernel<T>, rgrid, dim3(128), 0, stream, (const float*)p.part_o,
                    (const float*)p.part_lse, (T*)p.out, p.out_row_stride, p.num_heads, D,
                    p.num_splits);
    if (le != cudaSuccess)
      return fail(MVB200_ECUDA, "decode_reduce_kernel: %s", cudaGetErrorString(le));
    rc = check_launch("decode_reduce_kernel");
  }
  return rc;
}

template <typename T>
static int launch_generic(const void* q, int64_t q_row_stride, const void* k_cache,
                          const void* v_cache, const int32_t* block_tables, int64_t bt_row_stride,
                          const int32_t* context_lens, void* out, int64_t out_row_stride, int batch,
                          int num_heads, int num_kv_heads, int head_dim, int block_size,
                          int64_t num_blocks, int max_blocks, float scale, cudaStream_t stream) {
  dim3 grid(num_heads, batch);
  paged_decode_generic_kernel<T><<<grid, 128, 0, stream>>>(
      (const T*)q, q_row_stride, (const T*)k_cache, (const T*)v_cache, block_tables, bt_row_stride,
      context_lens, (T*)out, out_row_stride, num_heads, num_kv_heads, head_dim, block_size,
      num_blocks, max_blocks, scale * kLog2e);
  return check_launch("paged_decode_generic_kernel");
}

static int decode_stages() {
  static int v = 0;
  if (!v) {
    const char* e = getenv("MVB200_DECODE_STAGES");
    v = e ? atoi(e) : 3;
    if (v < 2) v = 2;
    if (v > 6) v = 6;
  }
  return v;
}

}  // namespace mvb200

extern "C" size_t mvb200_decode_workspace_bytes(int32_t batch, int32_t num_heads,
                                                int32_t head_dim, int32_t num_splits) {
  if (num_splits <= 1 || batch <= 0) return 0;
  return (size_t)batch * num_heads * num_splits * ((size_t)head_dim + 1) * sizeof(float);
}

extern "C" int32_t mvb200_decode_suggest_splits(int32_t batch, int32_t num_kv_heads,
                                                int32_t max_context, int32_t sm_count) {
  if (batch <= 0 || num_kv_heads <= 0) return 1;
  if (sm_count <= 0) sm_count = 148;
  const int64_t units = (int64_t)batch * num_kv_heads;
  const int tiles = max_context > 0 ? (max_context + 63) / 64 : 1;
  // Plenty of (sequence, kv head) units: two waves of two resident CTAs per SM.  Fewer units than SMs
  // (long-context, small-batch decode): one CTA per SM in a single wave — measured on B200 at
  // 1 x 32k x 4 kv heads: 128 CTAs 21 us, 296 CTAs 27 us, 512 CTAs 36 us (per-CTA prologue and the
  // split reduction grow with the split count).
  const int64_t want = units >= sm_count ? 4LL * sm_count : (int64_t)sm_count;
  int64_t s = units >= sm_count ? (want + units - 1) / units : want / units;
  const int max_by_work = tiles / 4 > 0 ? tiles / 4 : 1;  // keep >= 4 tiles (128 KiB) per CTA
  if (s > max_by_work) s = max_by_work;
  if (s < 1) s = 1;
  if (s > 128) s = 128;
  return (int32_t)s;
}

extern "C" int mvb200_paged_decode(const void* q, int64_t q_row_stride, const void* k_cache,
                                   const void* v_cache, const int32_t* block_tables,
                                   int64_t bt_row_stride, const int32_t* context_lens, void* out,
                                   int64_t out_row_stride, void* workspace, size_t workspace_bytes,
                                   int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                                   int32_t head_dim, int32_t block_size, int64_t num_blocks,
                                   int32_t max_blocks_per_seq, int32_t num_splits, float scale,
                                   int32_t dtype, void* stream) {
  using namespace mvb200;
  if (batch < 0 || num_heads <= 0 || num_kv_heads <= 0 || head_dim <= 0 || block_size <= 0)
    return fail(MVB200_EINVAL, "paged_decode: bad sizes");
  if (num_heads % num_kv_heads)
    return fail(MVB200_EINVAL, "paged_decode: num_heads %d not a multiple of num_kv_heads %d",
                num_heads, num_kv_heads);
  if (batch == 0) return MVB200_OK;
  if (!q || !out || !context_lens) return fail(MVB200_EINVAL, "paged_decode: null pointer");
  cudaStream_t s = (cudaStream_t)stream;
  const int eb = dtype == MVB200_F32 ? 4 : ((dtype == MVB200_F16 || dtype == MVB200_BF16) ? 2 : 0);
  if (!eb) return fail(MVB200_EINVAL, "paged_decode: unknown dtype %d", dtype);
  if (max_blocks_per_seq <= 0 || num_blocks <= 0) {
    // zero-width block table (engine warm-up): zeros, like page_attention.py:62-65
    for (int b = 0; b < batch; ++b) {
      cudaError_t e = cudaMemsetAsync((char*)out + (int64_t)b * out_row_stride * eb, 0,
                                      (size_t)num_heads * head_dim * eb, s);
      if (e != cudaSuccess) return fail(MVB200_ECUDA, "paged_decode: memset: %s", cudaGetErrorString(e));
    }
    note_kernel("decode_zero");
    return MVB200_OK;
  }
  if (!k_cache || !v_cache || !block_tables)
    return fail(MVB200_EINVAL, "paged_decode: null cache or block table");
  if (num_splits < 1) num_splits = 1;
  const int group = num_heads / num_kv_heads;
  const bool fast = eb == 2 && (head_dim == 64 || head_dim == 128) && block_size % 64 == 0 &&
                    group <= 16 && num_blocks < (1LL << 31) &&
                    ((uintptr_t)k_cache % 16 == 0) && ((uintptr_t)v_cache % 16 == 0) &&
                    ((uintptr_t)q % 4 == 0) && (q_row_stride % 2 == 0) && batch <= 65535 &&
                    num_splits <= 65535;
  if (fast) {
    DecodeParams p;
    p.q = q;
    p.q_row_stride = q_row_stride;
    p.block_tables = block_tables;
    p.bt_row_stride = bt_row_stride;
    p.context_lens = context_lens;
    p.out = out;
    p.out_row_stride = out_row_stride;
    p.num_heads = num_heads;
    p.num_kv_heads = num_kv_heads;
    p.group = group;
    p.max_blocks = max_blocks_per_seq;
    p.tiles_per_block = block_size / 64;
    p.num_blocks = (int32_t)num_blocks;
    p.num_splits = num_splits;
    p.stages = decode_stages();
    p.scale_log2 = scale * kLog2e;
    p.part_o = nullptr;
    p.part_lse = nullptr;
    if (num_splits > 1) {
      const size_t need = mvb200_decode_workspace_bytes(batch, num_heads, head_dim, num_splits);
      if (!workspace || workspace_bytes < need)
        return fail(MVB200_EINVAL, "paged_decode: workspace %zu B < %zu B needed for %d splits",
                    workspace_bytes, need, num_splits);
      p.part_o = (float*)workspace;
      p.part_lse = p.part_o + (size_t)batch * num_heads * num_splits * head_dim;
    }
    note_kernel("decode_tma_mma");
    if (dtype == MVB200_BF16) {
      return head_dim == 128 ? launch_fast<__nv_bfloat16, 128>(p, k_cache, v_cache, block_size, batch, s)
                             : launch_fast<__nv_bfloat16, 64>(p, k_cache, v_cache, block_size, batch, s);
    }
    return head_dim == 128 ? launch_fast<__half, 128>(p, k_cache, v_cache, block_size, batch, s)
                           : launch_fast<__half, 64>(p, k_cache, v_cache, block_size, batch, s);
  }
  if (head_dim > 256)
    return fail(MVB200_EUNSUPPORTED, "paged_decode: head_dim %d > 256", head_dim);
  if (batch > 65535) return fail(MVB200_EUNSUPPORTED, "paged_decode: batch %d > 65535", batch);
  note_kernel("decode_generic");
  switch (dtype) {
    case MVB200_F32:
      return launch_generic<float>(q, q_row_stride, k_cache, v_cache, block_tables, bt_row_stride,
                                   context_lens, out, out_row_stride, batch, num_heads,
                                   num_kv_heads, head_dim, block_size, num_blocks,
                                   max_blocks_per_seq, scale, s);
    case MVB200_F16:
      return launch_generic<__half>(q, q_row_stride, k_cache, v_cache, block_tables, bt_row_stride,
                                    context_lens, out, out_row_stride, batch, num_heads,
                                    num_kv_heads, head_dim, block_size, num_blocks,
                                    max_blocks_per_seq, scale, s);
    default:
      return launch_generic<__nv_bfloat16>(q, q_row_stride, k_cache, v_cache, block_tables,
                                           bt_row_stride, context_lens, out, out_row_stride, batch,
                                           num_heads, num_kv_heads, head_dim, block_size,
                                           num_blocks, max_blocks_per_seq, scale, s);
  }
}
