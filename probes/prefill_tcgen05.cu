// prefill_tcgen05.cu — varlen causal flash attention on the 5th-gen tensor cores (sm_100a).
//
// Replaces flash_attn_varlen_func(..., block_table=) at reference
// minivllm/models/layers/attention.py:422-433 (and its eager twin :467-504).
//
// One CTA = one 128-row query tile of one (sequence, head); 192 threads in three roles:
//   warp 0      TMA producer: Q tile once, then 64-key K and V tiles through a 2-stage ring.
//               Paged source: block-table entry -> coordinate of a 4-D tensor map over the cache
//               [NB, BS, KVH, D]; dense source: 3-D map over k/v [rows, KVH, D].
//   warp 1      MMA issuer (one lane): S_j = Q K_j^T (tcgen05.mma SS, M=128 N=64, 8 k-steps) into a
//               double-buffered TMEM accumulator, then O += P_{j-1} V_{j-1} (tcgen05.mma TS: A = bf16
//               P read from TMEM, B = V tile MN-major from smem, M=128 N=128, 4 k-steps).
//   warps 2..5  softmax warpgroup: thread r owns query row r — tcgen05.ld its 64 scores, bottom-right
//               causal mask on diagonal tiles, online softmax with exp2 and lazy (threshold 2^8)
//               rescale of the O accumulator in TMEM, bf16 P written back over S with tcgen05.st;
//               finally O / l -> global.
// TMEM: S0 [0,64) S1 [64,128) O [128,256) = 256 columns -> two CTAs per SM (smem 97 KB each), so one
// CTA's softmax overlaps the other's MMAs.  All hand-offs are mbarriers (TMA complete_tx, tcgen05.commit,
// thread arrivals); no __syncthreads in the main loop.
// Tensor-bound: 4*D flops per unmasked (q,k) pair per head.
#include <math.h>
#include <stdlib.h>

#include <type_traits>

#include "prefill_args.h"
#include "tcgen05.cuh"
#include "softmax_math.cuh"

namespace mvb200 {


namespace pf {

// Pipeline trace (debug builds only, -DPF_TRACE; see tools/prefill_trace.py): one CTA records clock64
// at every hand-off of its three roles so stalls can be attributed without guessing.
#ifdef PF_TRACE
__device__ unsigned long long g_trace[3 * 64 * 8];
__device__ unsigned long long g_cta[8192 * 4];  // per CTA: smid | n_kt << 32, t_start, t_first_S, t_end (globaltimer ns)
MVB_DEV unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)::"memory");
  return t;
}
#define PF_T(role, j, ev)                                                                     \
  do {                                                                                        \
    if (traced && (j) < 64) {                                                                 \
      unsigned long long t_;                                                                  \
      asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_)::"memory");                            \
      g_trace[((role) * 64 + (j)) * 8 + (ev)] = t_;                                           \
    }                                                                                         \
  } while (0)
#else
#define PF_T(role, j, ev) do {} while (0)
#endif

constexpr int kD = 128;
constexpr int kBM = 128;               // query rows per CTA
constexpr int kBN = 64;                // keys per tile
constexpr int kKVStages = 2;
constexpr int kThreads = 192;
constexpr int kQHalfBytes = kBM * 128;   // [128 rows][128 B]
constexpr int kQBytes = 2 * kQHalfBytes;
constexpr int kKVHalfBytes = kBN * 128;  // [64 rows][128 B]
constexpr int kKVBytes = 2 * kKVHalfBytes;
constexpr int kTmemCols = 256;
constexpr int kColS = 0, kColO = 128;
constexpr int kDefaultEmu = 16;
constexpr float kRescaleThreshold = 8.f;  // log2 domain: rescale O only when the max grew by > 2^8

// barrier slots
enum {
  BAR_Q_FULL = 0,
  BAR_K_FULL = 1,             // [kKVStages]
  BAR_K_EMPTY = 3,
  BAR_V_FULL = 5,
  BAR_V_EMPTY = 7,
  BAR_S_FULL = 9,             // [2]
  BAR_P_FULL = 11,            // [2]
  BAR_PV_DONE = 13,           // [2]
  BAR_COUNT = 15
};

struct Params {
  const int32_t* block_tables;
  int64_t bt_row_stride;
  const int32_t* cu_q;
  const int32_t* cu_k;
  void* out;
  int64_t out_row_stride;
  int32_t num_heads, num_kv_heads;
  int32_t tiles_per_block;  // block_size / 64 (paged)
  int32_t max_blocks;
  int32_t num_blocks;
  int32_t paged;
  float scale_log2;
};

template <typename T, int kEmu>
__global__ void __launch_bounds__(kThreads, 2)
prefill_tcgen05_kernel(const __grid_constant__ CUtensorMap tmap_q,
                       const __grid_constant__ CUtensorMap tmap_k,
                       const __grid_constant__ CUtensorMap tmap_v, const Params p) {
  constexpr bool kBf16 = sizeof(T) == 2 && !std::is_same<T, __half>::value;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t sQ = smem_base;
  const uint32_t sK = sQ + kQBytes;
  const uint32_t sV = sK + kKVStages * kKVBytes;
  const uint32_t bars = sV + kKVStages * kKVBytes;
  const uint32_t tmem_slot = bars + 8 * BAR_COUNT;
  auto bar = [&](int i) { return bars + 8u * i; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seq = blockIdx.z, head = blockIdx.y;
  const int q0 = p.cu_q[seq], lq = p.cu_q[seq + 1] - q0;
  const int k0 = p.cu_k[seq], lk = p.cu_k[seq + 1] - k0;
  const int n_qtiles = (lq + kBM - 1) / kBM;
  const int qt = (int)gridDim.x - 1 - (int)blockIdx.x;  // longest (last) query tiles first
  if (qt >= n_qtiles) return;
  const int row0 = qt * kBM;                       // first query row of this tile (within sequence)
  const int rows = min(kBM, lq - row0);
  const int shift = lk - lq;                       // query i sees keys 0..shift+i
  const int kv_end = min(lk, shift + row0 + rows);  // keys needed by the last valid row
  const int n_kt = kv_end > 0 ? (kv_end + kBN - 1) / kBN : 0;
  const int kvh = head / (p.num_heads / p.num_kv_heads);
  T* out = reinterpret_cast<T*>(p.out);
#ifdef PF_TRACE
  const int cta_lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  if (threadIdx.x == 64 && cta_lin < 8192) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_cta[cta_lin * 4 + 0] = (unsigned long long)smid | ((unsigned long long)n_kt << 32);
    g_cta[cta_lin * 4 + 1] = gtime();
  }
  const bool traced = blockIdx.x == 0 && blockIdx.y == gridDim.y / 2 && blockIdx.z == gridDim.z / 2 && lane == 0 &&
                      (warp <= 2);
#endif

  if (n_kt == 0) {  // Lk < Lq corner: rows that see no key produce zeros
    for (int i = threadIdx.x; i < rows * kD; i += blockDim.x)
      out[(int64_t)(q0 + row0 + i / kD) * p.out_row_stride + (int64_t)head * kD + i % kD] = from_f32<T>(0.f);
    return;
  }

  if (warp == 0) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  if (threadIdx.x == 32) {
    mbar_init(bar(BAR_Q_FULL), 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(bar(BAR_K_FULL + s), 1);
      mbar_init(bar(BAR_K_EMPTY + s), 1);
      mbar_init(bar(BAR_V_FULL + s), 1);
      mbar_init(bar(BAR_V_EMPTY + s), 1);
      mbar_init(bar(BAR_S_FULL + s), 1);
      mbar_init(bar(BAR_P_FULL + s), 4);  // one arrival per softmax warp
      mbar_init(bar(BAR_PV_DONE + s), 1);
    }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // =========================== TMA producer ===========================
    // The whole warp runs the loop (uniform control flow); one elected lane issues.
    if (elect_one()) {
      tma_prefetch_desc(&tmap_q);
      tma_prefetch_desc(&tmap_k);
      tma_prefetch_desc(&tmap_v);
      mbar_arrive_expect_tx(bar(BAR_Q_FULL), kQBytes);
      tma_load_3d(sQ, &tmap_q, bar(BAR_Q_FULL), 0, head, q0 + row0);
      tma_load_3d(sQ + kQHalfBytes, &tmap_q, bar(BAR_Q_FULL), 64, head, q0 + row0);
    }
    __syncwarp();
    const int32_t* bt = p.block_tables + (int64_t)seq * p.bt_row_stride;
    auto page_of = [&](int j) -> int {
      const int v = __ldg(bt + min(j / p.tiles_per_block, p.max_blocks - 1));
      return (v >= 0 && v < p.num_blocks) ? v : p.num_blocks;  // invalid -> OOB -> zero tile
    };
    int bid_next = p.paged ? page_of(0) : 0;
    for (int j = 0; j < n_kt; ++j) {
      const int s = j % kKVStages;
      const uint32_t ph = (j / kKVStages) & 1;
      const int bid = bid_next;
      if (p.paged && j + 1 < n_kt) bid_next = page_of(j + 1);
      const int sub = p.paged ? (j % p.tiles_per_block) * kBN : 0;
      PF_T(0, j, 0);
      mbar_wait(bar(BAR_K_EMPTY + s), ph ^ 1);
      PF_T(0, j, 1);
      const uint32_t kd = sK + s * kKVBytes;
      if (elect_one()) {
        mbar_arrive_expect_tx(bar(BAR_K_FULL + s), kKVBytes);
        if (p.paged) {
          tma_load_4d(kd, &tmap_k, bar(BAR_K_FULL + s), 0, kvh, sub, bid);
          tma_load_4d(kd + kKVHalfBytes, &tmap_k, bar(BAR_K_FULL + s), 64, kvh, sub, bid);
        } else {
          tma_load_3d(kd, &tmap_k, bar(BAR_K_FULL + s), 0, kvh, k0 + j * kBN);
          tma_load_3d(kd + kKVHalfBytes, &tmap_k, bar(BAR_K_FULL + s), 64, kvh, k0 + j * kBN);
        }
      }
      __syncwarp();
      mbar_wait(bar(BAR_V_EMPTY + s), ph ^ 1);
      PF_T(0, j, 2);
      const uint32_t vd = sV + s * kKVBytes;
      if (elect_one()) {
        mbar_arrive_expect_tx(bar(BAR_V_FULL + s), kKVBytes);
        if (p.paged) {
          tma_load_4d(vd, &tmap_v, bar(BAR_V_FULL + s), 0, kvh, sub, bid);
          tma_load_4d(vd + kKVHalfBytes, &tmap_v, bar(BAR_V_FULL + s), 64, kvh, sub, bid);
        } else {
          tma_load_3d(vd, &tmap_v, bar(BAR_V_FULL + s), 0, kvh, k0 + j * kBN);
          tma_load_3d(vd + kKVHalfBytes, &tmap_v, bar(BAR_V_FULL + s), 64, kvh, k0 + j * kBN);
        }
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    // Warp-uniform loop, one elected lane issues; descriptors advance by adding to their low word.
    constexpr uint32_t idesc_qk = umma_idesc(kBM, kBN, kBf16, false, false);
    constexpr uint32_t idesc_pv = umma_idesc(kBM, kD, kBf16, false, true);
    constexpr uint32_t hi_kmajor = umma_desc_hi(1024);  // Q, K: K-major, 8-row groups 1 KiB apart
    constexpr uint32_t hi_v = umma_desc_hi(1024);       // V: MN-major, 8-key groups 1 KiB apart
    const uint32_t q_lo = umma_desc_lo(sQ, 16);
    const uint32_t k_lo0 = umma_desc_lo(sK, 16);
    const uint32_t v_lo0 = umma_desc_lo(sV, kKVHalfBytes);  // LBO = distance between the 64-column halves
    const int last_valid = kv_end - (n_kt - 1) * kBN;  // keys in the last tile that exist
    auto issue_pv = [&](int i) {
      const int b = i & 1, s = i % kKVStages;
      mbar_wait(bar(BAR_V_FULL + s), (i / kKVStages) & 1);
      PF_T(1, i, 4);
      mbar_wait(bar(BAR_P_FULL + b), (i >> 1) & 1);
      PF_T(1, i, 5);
      tc_fence_after();
      // skip 16-key groups that lie entirely past the last needed key
      const int ksteps = (i == n_kt - 1) ? (last_valid + 15) / 16 : kBN / 16;
      if (elect_one()) {
        const uint32_t v_lo = v_lo0 + s * (kKVBytes >> 4);
        const uint32_t p_col = tmem + kColS + b * kBN;
#pragma unroll
        for (int kk = 0; kk < kBN / 16; ++kk) {
          if (kk < ksteps)
            umma_ts(tmem + kColO, p_col + kk * 8, umma_desc(hi_v, v_lo + kk * (16 * 128 >> 4)), idesc_pv,
                    (i > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(bar(BAR_V_EMPTY + s));
        umma_commit(bar(BAR_PV_DONE + b));
      }
      __syncwarp();
      PF_T(1, i, 6);
    };
    mbar_wait(bar(BAR_Q_FULL), 0);
    for (int j = 0; j < n_kt; ++j) {
      const int b = j & 1, s = j % kKVStages;
      PF_T(1, j, 0);
      mbar_wait(bar(BAR_K_FULL + s), (j / kKVStages) & 1);
      PF_T(1, j, 1);
      if (j >= 2) mbar_wait(bar(BAR_PV_DONE + b), ((j - 2) >> 1) & 1);  // S/P buffer b is free again
      PF_T(1, j, 2);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t k_lo = k_lo0 + s * (kKVBytes >> 4);
        const uint32_t s_col = tmem + kColS + b * kBN;
#pragma unroll
        for (int kk = 0; kk < kD / 16; ++kk) {
          // 16 elements = 32 B inside the 128-B swizzle row; the next 64 elements are the next half-tile
          const uint32_t qoff = ((kk >> 2) * kQHalfBytes + (kk & 3) * 32) >> 4;
          const uint32_t koff = ((kk >> 2) * kKVHalfBytes + (kk & 3) * 32) >> 4;
          umma_ss(s_col, umma_desc(hi_kmajor, q_lo + qoff), umma_desc(hi_kmajor, k_lo + koff), idesc_qk,
                  kk > 0 ? 1u : 0u);
        }
        umma_commit(bar(BAR_K_EMPTY + s));
        umma_commit(bar(BAR_S_FULL + b));
      }
      __syncwarp();
      PF_T(1, j, 3);
      if (j >= 1) issue_pv(j - 1);
    }
    issue_pv(n_kt - 1);
  } else {
    // =========================== softmax warpgroup ===========================
    const int quarter = warp & 3;                 // TMEM lanes this warp may touch
    const int r = quarter * 32 + lane;            // query row within the tile
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const int q_limit = shift + row0 + r;         // largest key index this row may see
    float m_ref = -INFINITY, l_run = 0.f;

    for (int j = 0; j < n_kt; ++j) {
      const int b = j & 1;
      PF_T(2, j, 0);
      mbar_wait(bar(BAR_S_FULL + b), (j >> 1) & 1);
      PF_T(2, j, 1);
#ifdef PF_TRACE
      if (j == 0 && threadIdx.x == 64 && cta_lin < 8192) g_cta[cta_lin * 4 + 2] = gtime();
#endif
      tc_fence_after();
      float s[kBN];
      {
        uint32_t raw0[32], raw1[32];
        tmem_ld32(lane_base + kColS + b * kBN, raw0);
        tmem_ld32(lane_base + kColS + b * kBN + 32, raw1);
        tmem_wait_ld();
        PF_T(2, j, 2);
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          s[c] = __uint_as_float(raw0[c]);
          s[32 + c] = __uint_as_float(raw1[c]);
        }
      }
      const int col0 = j * kBN;
      if (col0 + kBN - 1 > shift + row0) {  // tile touches the diagonal (CTA-uniform test)
        const int lim = q_limit - col0;
#pragma unroll
        for (int c = 0; c < kBN; ++c) s[c] = c <= lim ? s[c] : -INFINITY;
      }
      float mx = fmaxf(s[0], s[1]);
#pragma unroll
      for (int c = 2; c < kBN; c += 2) mx = fmaxf(mx, fmaxf(s[c], s[c + 1]));
      mx *= p.scale_log2;
      // lazy rescale: keep exponentiating against m_ref until the true max outruns it by 2^8
      const bool grow = mx > m_ref + kRescaleThreshold;
      if (__any_sync(0xffffffffu, grow)) {
        const float m_new = grow ? mx : m_ref;
        if (j > 0) {
          const float sc = ex2_mufu(m_ref - m_new);  // 1 for lanes that keep their reference
          l_run *= sc;
          mbar_wait(bar(BAR_PV_DONE + ((j - 1) & 1)), ((j - 1) >> 1) & 1);  // O is quiescent
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < kD; c += 32) {
            uint32_t o[32];
            tmem_ld32(lane_base + kColO + c, o);
            tmem_wait_ld();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * sc);
            tmem_st32(lane_base + kColO + c, o);
          }
        }
        m_ref = m_new;
      }
      const float m_use = m_ref == -INFINITY ? 0.f : m_ref;
      PF_T(2, j, 3);
      uint32_t pk[kBN / 2];
      const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);
      const uint64_t negm2 = pack2(-m_use, -m_use);
      uint64_t sum2 = pack2(0.f, 0.f);
#pragma unroll
      for (int c = 0; c < kBN; c += 2) {
        const uint64_t x = fma2(pack2(s[c], s[c + 1]), scale2, negm2);
        float p0, p1;
        // kEmu of every 64 columns take the FMA-pipe exponential, interleaved with the MUFU ones
        if (kEmu > 0 && ((c / 2) * (kEmu / 2)) % (kBN / 2) < kEmu / 2) {
          ex2_poly2(x, p0, p1);
        } else {
          float x0, x1;
          unpack2(x, x0, x1);
          p0 = ex2_mufu(x0);
          p1 = ex2_mufu(x1);
        }
        sum2 = add2(sum2, pack2(p0, p1));
        pk[c / 2] = Mma16816<T>::pack(p0, p1);
      }
      float sum, sum_hi;
      unpack2(sum2, sum, sum_hi);
      sum += sum_hi;
      l_run += sum;
      PF_T(2, j, 4);
      tmem_st32(lane_base + kColS + b * kBN, pk);  // P overwrites the first half of S_b
      tmem_wait_st();
      PF_T(2, j, 5);

      if (j == n_kt - 1 && warp == 2) {
        // keys >= kv_end in the last tile were never written: their V rows may hold NaN/Inf bit
        // patterns, and 0 * NaN would poison O.  Zero the rows of the last (partially used)
        // 16-key group before the MMA reads them (whole groups past it are skipped by the issuer).
        const int valid = kv_end - col0;
        const int stop = min(kBN, (valid + 15) & ~15);
        if (valid < stop) {
          const int st = j % kKVStages;
          mbar_wait(bar(BAR_V_FULL + st), (j / kKVStages) & 1);
          uint8_t* vg = smem_gen + (sV - smem_base) + st * kKVBytes;
          for (int row = valid; row < stop; ++row)
            if (lane < 16)
              *reinterpret_cast<uint4*>(vg + (lane >> 3) * kKVHalfBytes + row * 128 + (lane & 7) * 16) =
                  make_uint4(0, 0, 0, 0);
          fence_proxy_async_smem();
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(BAR_P_FULL + b));
      PF_T(2, j, 6);
    }

    // ---- epilogue: O / l -> global ----
    mbar_wait(bar(BAR_PV_DONE + ((n_kt - 1) & 1)), ((n_kt - 1) >> 1) & 1);
    tc_fence_after();
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    T* orow = out + (int64_t)(q0 + row0 + r) * p.out_row_stride + (int64_t)head * kD;
#pragma unroll
    for (int c = 0; c < kD; c += 32) {
      uint32_t o[32];
      tmem_ld32(lane_base + kColO + c, o);
      tmem_wait_ld();
      if (r < rows) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 w;
          w.x = Mma16816<T>::pack(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
          w.y = Mma16816<T>::pack(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
          w.z = Mma16816<T>::pack(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
          w.w = Mma16816<T>::pack(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + c + i) = w;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kTmemCols);
#ifdef PF_TRACE
  if (threadIdx.x == 64 && cta_lin < 8192) g_cta[cta_lin * 4 + 3] = gtime();
#endif
}

template <typename T, int kEmu>
static int launch(const PrefillArgs& a, cudaStream_t stream) {
  CUtensorMap mq, mk, mv;
  {
    const uint64_t dims[3] = {(uint64_t)kD, (uint64_t)a.num_heads, (uint64_t)a.total_q};
    const uint64_t str[3] = {0, (uint64_t)kD * 2, (uint64_t)a.q_row_stride * 2};
    const uint32_t box[3] = {64, 1, (uint32_t)kBM};
    int rc = make_tensor_map(&mq, a.q, 2, 3, dims, str, box, true);
    if (rc) return rc;
  }
  const bool paged = a.block_tables != nullptr;
  if (paged) {
    const uint64_t dims[4] = {(uint64_t)kD, (uint64_t)a.num_kv_heads, (uint64_t)a.block_size,
                              (uint64_t)a.num_blocks};
    const uint64_t str[4] = {0, (uint64_t)kD * 2, (uint64_t)a.num_kv_heads * kD * 2,
                             (uint64_t)a.block_size * a.num_kv_heads * kD * 2};
    const uint32_t box[4] = {64, 1, (uint32_t)kBN, 1};
    int rc = make_tensor_map(&mk, a.k_cache, 2, 4, dims, str, box, true);
    if (rc) return rc;
    rc = make_tensor_map(&mv, a.v_cache, 2, 4, dims, str, box, true);
    if (rc) return rc;
  } else {
    const uint64_t dims[3] = {(uint64_t)kD, (uint64_t)a.num_kv_heads, (uint64_t)a.num_kv_rows};
    const uint32_t box[3] = {64, 1, (uint32_t)kBN};
    const uint64_t strk[3] = {0, (uint64_t)kD * 2, (uint64_t)a.k_row_stride * 2};
    const uint64_t strv[3] = {0, (uint64_t)kD * 2, (uint64_t)a.v_row_stride * 2};
    int rc = make_tensor_map(&mk, a.k, 2, 3, dims, strk, box, true);
    if (rc) return rc;
    rc = make_tensor_map(&mv, a.v, 2, 3, dims, strv, box, true);
    if (rc) return rc;
  }
  Params p;
  p.block_tables = a.block_tables;
  p.bt_row_stride = a.bt_row_stride;
  p.cu_q = a.cu_q;
  p.cu_k = a.cu_k;
  p.out = a.out;
  p.out_row_stride = a.out_row_stride;
  p.num_heads = a.num_heads;
  p.num_kv_heads = a.num_kv_heads;
  p.tiles_per_block = paged ? a.block_size / kBN : 1;
  p.max_blocks = paged ? a.max_blocks : 1;
  p.num_blocks = paged ? (int32_t)a.num_blocks : 0;
  p.paged = paged ? 1 : 0;
  p.scale_log2 = a.scale * kLog2e;

  constexpr int smem = kQBytes + 2 * kKVStages * kKVBytes + 8 * BAR_COUNT + 16 + 1024;
  auto kern = prefill_tcgen05_kernel<T, kEmu>;
  static SmemOptIn optin;
  if (int rc = optin_dynamic_smem((const void*)kern, smem, &optin, "prefill_tcgen05")) return rc;
  const int qtiles = (a.max_seqlen_q + kBM - 1) / kBM;
  dim3 grid(qtiles, a.num_heads, a.num_seqs);
  kern<<<grid, kThreads, smem, stream>>>(mq, mk, mv, p);
  return check_launch("prefill_tcgen05_kernel");
}

}  // namespace pf

// returns -1 when the request is outside what the tcgen05 kernel covers
int prefill_tcgen05_dispatch(const PrefillArgs& a, cudaStream_t stream) {
  if (a.dtype != MVB200_BF16 && a.dtype != MVB200_F16) return -1;
  if (a.head_dim != pf::kD) return -1;
  if (a.max_seqlen_q <= 0 || a.num_heads > 65535 || a.num_seqs > 65535) return -1;
  auto al16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  if (!al16(a.q) || !al16(a.out) || (a.q_row_stride % 8) || (a.out_row_stride % 8)) return -1;
  if (a.block_tables) {
    if (a.block_size % pf::kBN || !al16(a.k_cache) || !al16(a.v_cache) ||
        a.num_blocks >= (1LL << 31))
      return -1;
  } else {
    if (!al16(a.k) || !al16(a.v) || (a.k_row_stride % 8) || (a.v_row_stride % 8) ||
        a.num_kv_rows <= 0)
      return -1;
  }
  note_kernel("prefill_tcgen05");
  // Columns (of every 64) whose exponential runs on the FMA pipe instead of MUFU.  Tuning knob for
  // profiling only (MVB200_PREFILL_EMU = 0 | 8 | 16 | 24); the default is the measured optimum.
  static const int emu = [] {
    const char* e = getenv("MVB200_PREFILL_EMU");
    const int v = e ? atoi(e) : pf::kDefaultEmu;
    return (v == 0 || v == 8 || v == 16 || v == 24) ? v : pf::kDefaultEmu;
  }();
  const bool bf = a.dtype == MVB200_BF16;
  switch (emu) {
    case 0: return bf ? pf::launch<__nv_bfloat16, 0>(a, stream) : pf::launch<__half, 0>(a, stream);
    case 8: return bf ? pf::launch<__nv_bfloat16, 8>(a, stream) : pf::launch<__half, 8>(a, stream);
    case 24: return bf ? pf::launch<__nv_bfloat16, 24>(a, stream) : pf::launch<__half, 24>(a, stream);
    default: return bf ? pf::launch<__nv_bfloat16, 16>(a, stream) : pf::launch<__half, 16>(a, stream);
  }
}

}  // namespace mvb200

#ifdef PF_TRACE
extern "C" __attribute__((visibility("default"))) int mvb200_debug_read_prefill_trace(unsigned long long* dst) {
  return (int)cudaMemcpyFromSymbol(dst, mvb200::pf::g_trace, sizeof(mvb200::pf::g_trace));
}
extern "C" __attribute__((visibility("default"))) int mvb200_debug_read_prefill_ctas(unsigned long long* dst) {
  return (int)cudaMemcpyFromSymbol(dst, mvb200::pf::g_cta, sizeof(mvb200::pf::g_cta));
}
#endif
