// prefill_dbs.cu — persistent varlen causal flash attention with DOUBLE-BUFFERED S (sm_100a).
//
// Replaces flash_attn_varlen_func(..., block_table=) at reference
// minivllm/models/layers/attention.py:422-433 (and its eager twin :467-504).
//
// Third generation of the prefill kernel.  prefill_fa2q.cu (two 128-row query tiles per CTA, 128-key tiles,
// bf16 P aliasing S in TMEM) runs each query tile as a serial chain
//     S(j) -> softmax -> P(j) V -> S(j+1)          (S(j+1) overwrites P(j))
// whose period is softmax + 1024 tensor-pipe clocks: 3500 clocks per 128 keys against 2048 of MMA work
// (profiles/r01_prefill_v2a_trace_emu32.log), because a lone softmax warp per SM sub-partition issues at
// 2-3 clocks per instruction (profiles/r02_pipe_rates.txt) and sits on the critical path.  Splitting the
// softmax of a tile over two warpgroups did not help (profiles/r02_prefill_ab_run5.jsonl).
// This kernel takes the softmax OFF the critical path instead:
//   * 64-key steps: S is [128 x 64] fp32 = 64 TMEM columns, so each query tile owns TWO S buffers
//     (2 tiles x (2 x 64 S + 128 O) = 512 columns).  S(j+1) is computed while the warpgroup still works on
//     S(j); P(j) overwrites the first 32 columns of S(j)'s buffer and S(j+2) reuses it after P(j) V.
//   * per query tile the issue order is  QK(0) QK(1) | P(0)V QK(2) | P(1)V QK(3) | ...: tcgen05.mma
//     executes in issue order, so QK(j+2) cannot overtake the P(j) V that still reads the buffer.
//   * both softmax warpgroups are busy all the time (two warps per sub-partition: better issue rates), the
//     tensor pipe alternates between the two chains.
// Cost: S = Q K^T with N = 64 runs at 2/3 rate (192 B/clk of shared-memory operand reads against a
// 128 B/clk limit, probes/umma_pacing), i.e. 640 instead of 512 pipe clocks per 64 keys and tile.
// Roles (384 threads): warp 0 TMA producer, warps 1 / 2 MMA issuers of query tiles 0 / 1, warp 3 idle,
// warps 4-7 / 8-11 the softmax warpgroups of query tiles 0 / 1 (thread = query row).
// Tensor-bound: 4*D flops per unmasked (q,k) pair per head.
#include <math.h>
#include <stdlib.h>

#include <type_traits>

#include "prefill_args.h"
#include "softmax_math.cuh"
#include "tcgen05.cuh"

namespace mvb200 {
namespace dbs {

using pf::add2;
using pf::ex2_mufu;
using pf::ex2_poly2;
using pf::fma2;
using pf::pack2;
using pf::unpack2;

constexpr int kBM = 128;                  // rows per query tile
constexpr int kQTiles = 2;                // query tiles (chains) per work item
constexpr int kBN = 64;                   // keys per step (= one TMA box = kvcache_block_size granularity)
constexpr int kStages = 4;                // K ring depth = V ring depth
constexpr int kThreads = 384;
constexpr int kQHalfBytes = kBM * 128;    // [128 rows][128 B] = one 64-column slab of a Q tile (head_dim / 64 slabs)
constexpr int kKVHalfBytes = kBN * 128;   // [64 keys][128 B]
constexpr int kTmemCols = 512;
constexpr int kColS = 0;                  // + t * 128 + buffer * 64
constexpr int kColO = 256;                // + t * 128
constexpr int kMaxLevels = 512;           // query blocks per sequence (256 rows each): 131072 tokens
constexpr int kDefaultEmu = 16;           // of every 64 columns, exponentials done on the FMA pipe
constexpr float kRescaleThreshold = 8.f;  // log2 domain: rescale O only when the max grew by > 2^8

enum {
  BAR_Q_FULL = 0,                       // [2]  TMA -> MMA
  BAR_Q_EMPTY = 2,                      // [2]  MMA (commit after the item's last Q_t K^T) -> TMA
  BAR_K_FULL = 4,                       // [kStages]
  BAR_K_EMPTY = BAR_K_FULL + kStages,   // [kStages]  MMA (commit after the last Q_t K^T of the step) -> TMA
  BAR_V_FULL = BAR_K_EMPTY + kStages,
  BAR_V_EMPTY = BAR_V_FULL + kStages,
  BAR_S_FULL = BAR_V_EMPTY + kStages,   // [2 tiles][2 buffers]  MMA -> softmax warpgroup t
  BAR_P_FULL = BAR_S_FULL + 4,          // [2 tiles][2 buffers]  softmax warpgroup t (4 warp arrivals) -> MMA
  BAR_O_DONE = BAR_P_FULL + 4,          // [2]  MMA (commit after every P_t V) -> softmax warpgroup t (lazy rescale)
  BAR_O_FINAL = BAR_O_DONE + 2,         // [2]  MMA (commit after the item's last P_t V) -> softmax warpgroup t (epilogue)
  BAR_COUNT = BAR_O_FINAL + 2
};

struct Params {
  const int32_t* block_tables;
  int64_t bt_row_stride;
  const int32_t* cu_q;
  const int32_t* cu_k;
  void* out;
  int64_t out_row_stride;
  int32_t num_heads, num_kv_heads, num_seqs;
  int32_t pages_per_block;  // block_size / 64 (paged)
  int32_t max_blocks;
  int32_t num_blocks;
  int32_t paged;
  int32_t levels;           // ceil(max_seqlen_q / 256)
  int32_t max_seqlen_q;
  float scale_log2;
};

// One work item, decoded identically by every role from its linear index.
struct Item {
  int seq, head, kvh;
  int q0, k0, lk;      // first packed query / key row of the sequence, key count
  int lq;              // query rows of the sequence
  int shift;           // query row i sees keys 0..shift+i
  int row0[kQTiles];   // first query row of tile t within the sequence
  int rows[kQTiles];   // valid rows of tile t (0: tile absent)
  int kv_end[kQTiles]; // keys tile t needs
  int n[kQTiles];      // 64-key steps tile t needs (0: nothing to attend to)
  int n_kt;            // K/V tiles the item streams = max(n[0], n[1])
  bool valid;
};

// Items are ordered by query-block index, highest first: within one launch that is longest-first
// (a later query block attends to more keys), which is what a static round-robin needs to balance.
MVB_DEV Item decode_item(const Params& p, int w) {
  Item it;
  const int per_level = p.num_seqs * p.num_heads;
  const int level = p.levels - 1 - w / per_level;
  const int rem = w % per_level;
  it.seq = rem / p.num_heads;
  it.head = rem % p.num_heads;
  it.kvh = it.head / (p.num_heads / p.num_kv_heads);
  it.q0 = __ldg(p.cu_q + it.seq);
  const int lq = __ldg(p.cu_q + it.seq + 1) - it.q0;
  it.k0 = __ldg(p.cu_k + it.seq);
  it.lk = __ldg(p.cu_k + it.seq + 1) - it.k0;
  it.lq = lq;
  it.shift = it.lk - lq;
  it.n_kt = 0;
#pragma unroll
  for (int t = 0; t < kQTiles; ++t) {
    it.row0[t] = level * (kQTiles * kBM) + t * kBM;
    it.rows[t] = max(0, min(kBM, lq - it.row0[t]));
    it.kv_end[t] = it.rows[t] > 0 ? max(0, min(it.lk, it.shift + it.row0[t] + it.rows[t])) : 0;
    it.n[t] = (it.kv_end[t] + kBN - 1) / kBN;
    it.n_kt = max(it.n_kt, it.n[t]);
  }
  it.valid = it.rows[0] > 0;
  return it;
}

template <typename T, int kEmu, int kD>
__global__ void __launch_bounds__(kThreads, 1)
prefill_dbs_kernel(const __grid_constant__ CUtensorMap tmap_q,
                   const __grid_constant__ CUtensorMap tmap_k,
                   const __grid_constant__ CUtensorMap tmap_v, const Params p) {
  static_assert(kD == 64 || kD == 128, "head_dim 64 or 128");
  constexpr bool kBf16 = sizeof(T) == 2 && !std::is_same<T, __half>::value;
  constexpr int kSlabs = kD / 64;                       // 128-byte column slabs per row
  constexpr int kQTileBytes = kSlabs * kQHalfBytes;
  constexpr int kKVTileBytes = kSlabs * kKVHalfBytes;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t sQ = smem_base;                          // [2 tiles][32 KB]
  const uint32_t sK = sQ + kQTiles * kQTileBytes;         // [kStages][16 KB]
  const uint32_t sV = sK + kStages * kKVTileBytes;
  const uint32_t bars = sV + kStages * kKVTileBytes;
  const uint32_t tmem_slot = bars + 8 * BAR_COUNT;
  auto bar = [&](int i) { return bars + 8u * i; };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_items = p.levels * p.num_seqs * p.num_heads;

  if (warp == 0) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  if (threadIdx.x == 32) {
    for (int i = 0; i < BAR_COUNT; ++i)
      mbar_init(bar(i), (i >= BAR_P_FULL && i < BAR_P_FULL + 4) ? 4
                        : ((i >= BAR_K_EMPTY && i < BAR_K_EMPTY + kStages) || (i >= BAR_V_EMPTY && i < BAR_V_EMPTY + kStages)) ? 2 : 1);
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (elect_one()) {
      tma_prefetch_desc(&tmap_q);
      tma_prefetch_desc(&tmap_k);
      tma_prefetch_desc(&tmap_v);
    }
    __syncwarp();
    uint32_t q_uses[kQTiles] = {0, 0};  // loads issued so far into Q buffer t
    uint32_t kv_cnt = 0;                // K/V tiles issued so far (stage = cnt % kStages)
    for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
      const Item it = decode_item(p, w);
      if (it.lq > p.levels * kQTiles * kBM) {
        // the caller's max_seqlen_q understates a sequence: its tail rows would silently stay unwritten
        if (lane == 0) printf("mvb200: prefill: max_seqlen_q %d < query length %d of sequence %d\n",
                              p.max_seqlen_q, it.lq, it.seq);
        __trap();
      }
      if (!it.valid || it.n_kt == 0) continue;
#pragma unroll
      for (int t = 0; t < kQTiles; ++t) {
        if (it.n[t] == 0) continue;
        mbar_wait(bar(BAR_Q_EMPTY + t), (q_uses[t] & 1) ^ 1);  // previous item's Q K^T are done
        ++q_uses[t];
        if (elect_one()) {
          mbar_arrive_expect_tx(bar(BAR_Q_FULL + t), kQTileBytes);
          const uint32_t dst = sQ + t * kQTileBytes;
#pragma unroll
          for (int h = 0; h < kSlabs; ++h)
            tma_load_3d(dst + h * kQHalfBytes, &tmap_q, bar(BAR_Q_FULL + t), 64 * h, it.head, it.q0 + it.row0[t]);
        }
        __syncwarp();
      }
      const int32_t* bt = p.block_tables + (int64_t)it.seq * p.bt_row_stride;
      for (int j = 0; j < it.n_kt; ++j, ++kv_cnt) {
        const int s = kv_cnt % kStages;
        const uint32_t ph = (kv_cnt / kStages) & 1;
        // the 64-key page of this step: block id (paged) or packed row (dense)
        int c_blk, c_sub = 0;
        if (p.paged) {
          const int v = __ldg(bt + min(j / p.pages_per_block, p.max_blocks - 1));
          c_blk = (v >= 0 && v < p.num_blocks) ? v : p.num_blocks;  // invalid -> OOB -> zero fill
          c_sub = (j % p.pages_per_block) * kBN;
        } else {
          c_blk = it.k0 + j * kBN;
        }
        mbar_wait(bar(BAR_K_EMPTY + s), ph ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(bar(BAR_K_FULL + s), kKVTileBytes);
          const uint32_t dst = sK + s * kKVTileBytes;
#pragma unroll
          for (int h = 0; h < kSlabs; ++h) {
            if (p.paged) tma_load_4d(dst + h * kKVHalfBytes, &tmap_k, bar(BAR_K_FULL + s), 64 * h, it.kvh, c_sub, c_blk);
            else tma_load_3d(dst + h * kKVHalfBytes, &tmap_k, bar(BAR_K_FULL + s), 64 * h, it.kvh, c_blk);
          }
        }
        __syncwarp();
        mbar_wait(bar(BAR_V_EMPTY + s), ph ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(bar(BAR_V_FULL + s), kKVTileBytes);
          const uint32_t dst = sV + s * kKVTileBytes;
#pragma unroll
          for (int h = 0; h < kSlabs; ++h) {
            if (p.paged) tma_load_4d(dst + h * kKVHalfBytes, &tmap_v, bar(BAR_V_FULL + s), 64 * h, it.kvh, c_sub, c_blk);
            else tma_load_3d(dst + h * kKVHalfBytes, &tmap_v, bar(BAR_V_FULL + s), 64 * h, it.kvh, c_blk);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1 || warp == 2) {
    // =========================== MMA issuers: warp 1 drives query tile 0, warp 2 query tile 1 ===========================
    // Per query tile the chain is  QK(0) QK(1) | wait P(0): P(0)V, QK(2) | wait P(1): P(1)V, QK(3) | ...
    // (step x of a chain, counted across items, uses S buffer x & 1).  tcgen05.mma from one thread executes in
    // issue order, which is what keeps QK(j+2) behind the P(j)V that still reads its buffer; the two chains have
    // nothing to order between them, so each gets its own issuing thread: neither waits on the other's P, and
    // the 60-clock issue-to-issue latency of an N = 64 SS MMA (probes/umma_pacing: 60.6 clocks from one issuer,
    // 48 from two) overlaps with the other chain's work.  A K/V stage is released by two arrivals, one per
    // chain: a tcgen05.commit where the chain read the tile, a plain arrive (after the tile has landed, i.e. in
    // the right barrier phase) where it did not.
    const int t = warp - 1;
    constexpr uint32_t idesc_qk = umma_idesc(kBM, kBN, kBf16, false, false);
    constexpr uint32_t idesc_pv = umma_idesc(kBM, kD, kBf16, false, true);
    constexpr uint32_t desc_hi = umma_desc_hi(1024);  // 8-row (K-major) / 8-key (MN-major) groups 1 KiB apart
    const uint32_t q_lo = umma_desc_lo(sQ, 16) + t * (kQTileBytes >> 4);
    const uint32_t k_lo0 = umma_desc_lo(sK, 16);
    const uint32_t v_lo0 = umma_desc_lo(sV, kKVHalfBytes);  // LBO = distance between the 64-column halves
    const uint32_t s_base = tmem + kColS + t * 128;
    const uint32_t o_col = tmem + kColO + t * 128;
    uint32_t q_uses = 0;
    uint32_t gq = 0;  // Q K^T issued so far (= step index of the next one)
    uint32_t gp = 0;  // P V issued so far
    uint32_t kv_cnt = 0;

    for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
      const Item it = decode_item(p, w);
      if (!it.valid || it.n_kt == 0) continue;
      const int n_t = t ? it.n[1] : it.n[0];
      const int kv_end_t = t ? it.kv_end[1] : it.kv_end[0];
      // S(jj) = Q K(jj)^T : 8 k-steps of 16 elements (32 B inside the 128-B swizzle row, halves apart)
      auto issue_qk = [&](int jj) {
        const uint32_t cnt = kv_cnt + jj;
        const int s = cnt % kStages;
        mbar_wait(bar(BAR_K_FULL + s), (cnt / kStages) & 1);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t k_lo = k_lo0 + s * (kKVTileBytes >> 4);
          const uint32_t s_col = s_base + (gq & 1) * kBN;
#pragma unroll
          for (int kk = 0; kk < kD / 16; ++kk) {
            const uint32_t qoff = ((kk >> 2) * kQHalfBytes + (kk & 3) * 32) >> 4;
            const uint32_t koff = ((kk >> 2) * kKVHalfBytes + (kk & 3) * 32) >> 4;
            umma_ss(s_col, umma_desc(desc_hi, q_lo + qoff), umma_desc(desc_hi, k_lo + koff), idesc_qk, kk > 0 ? 1u : 0u);
          }
          umma_commit(bar(BAR_S_FULL + t * 2 + (gq & 1)));
          umma_commit(bar(BAR_K_EMPTY + s));
          if (jj == n_t - 1) umma_commit(bar(BAR_Q_EMPTY + t));
        }
        __syncwarp();
        ++gq;
      };

      if (n_t > 0) {
        mbar_wait(bar(BAR_Q_FULL + t), q_uses & 1);
        ++q_uses;
        issue_qk(0);
        if (n_t > 1) issue_qk(1);
      }
      for (int j = 0; j < n_t; ++j) {
        const uint32_t cnt = kv_cnt + j;
        const int s = cnt % kStages;
        const int ksteps = j + 1 < n_t ? kBN / 16 : (kv_end_t - j * kBN + 15) / 16;
        mbar_wait(bar(BAR_V_FULL + s), (cnt / kStages) & 1);
        mbar_wait(bar(BAR_P_FULL + t * 2 + (gp & 1)), (gp >> 1) & 1);
        tc_fence_after();
        if (elect_one()) {
          // A = bf16 P in TMEM (8 columns per 16 keys), B = V MN-major, 16 keys = 2 KiB
          const uint32_t v_lo = v_lo0 + s * (kKVTileBytes >> 4);
          const uint32_t p_col = s_base + (gp & 1) * kBN;
#pragma unroll
          for (int kk = 0; kk < kBN / 16; ++kk) {
            if (kk < ksteps)
              umma_ts(o_col, p_col + kk * 8, umma_desc(desc_hi, v_lo + kk * (16 * 128 >> 4)), idesc_pv,
                      (j > 0 || kk > 0) ? 1u : 0u);
          }
          umma_commit(bar(BAR_O_DONE + t));
          umma_commit(bar(BAR_V_EMPTY + s));
          if (j == n_t - 1) umma_commit(bar(BAR_O_FINAL + t));
        }
        __syncwarp();
        ++gp;
        if (j + 2 < n_t) issue_qk(j + 2);
      }
      // tiles only the other chain reads: hand this chain's share of the release over once they have landed
      for (int j = n_t; j < it.n_kt; ++j) {
        const uint32_t cnt = kv_cnt + j;
        const int s = cnt % kStages;
        mbar_wait(bar(BAR_K_FULL + s), (cnt / kStages) & 1);
        if (lane == 0) mbar_arrive(bar(BAR_K_EMPTY + s));
        mbar_wait(bar(BAR_V_FULL + s), (cnt / kStages) & 1);
        if (lane == 0) mbar_arrive(bar(BAR_V_EMPTY + s));
        __syncwarp();
      }
      kv_cnt += it.n_kt;
    }
  } else if (warp < 4) {
    // idle warps: keep the softmax warpgroups aligned to TMEM lane quarters (warp % 4)
  } else {
    // =========================== softmax warpgroups (thread = query row) ===========================
    const int t = (warp - 4) >> 2;                // query tile (chain) of this warpgroup
    const int quarter = warp & 3;                 // TMEM lanes this warp may touch
    const int r = quarter * 32 + lane;            // query row within the tile
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t s_col0 = lane_base + kColS + t * 128;
    const uint32_t o_col = lane_base + kColO + t * 128;
    T* out = reinterpret_cast<T*>(p.out);
    uint32_t g = 0;        // steps consumed by this warpgroup so far (S buffer g & 1; = P V commits it may wait for)
    uint32_t items_done = 0;  // items whose O this warpgroup has written out (phase of BAR_O_FINAL)
    uint32_t kv_cnt = 0;   // K/V tiles streamed by the CTA before the current item

    for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
      const Item it = decode_item(p, w);
      if (!it.valid) continue;
      const int kv_base = kv_cnt;
      kv_cnt += it.n_kt;
      // scalar copies (t is a run-time value: indexing the arrays with it would push Item into local memory)
      const int rows_t = t ? it.rows[1] : it.rows[0];
      const int row0_t = t ? it.row0[1] : it.row0[0];
      const int kv_end_t = t ? it.kv_end[1] : it.kv_end[0];
      const int n_t = t ? it.n[1] : it.n[0];
      if (rows_t == 0) continue;
      T* orow = out + (int64_t)(it.q0 + row0_t + r) * p.out_row_stride + (int64_t)it.head * kD;
      if (n_t == 0) {  // Lk < Lq corner: rows that see no key produce zeros
        if (r < rows_t)
#pragma unroll
          for (int i = 0; i < kD; i += 8) *reinterpret_cast<uint4*>(orow + i) = make_uint4(0, 0, 0, 0);
        continue;
      }
      const int q_limit = it.shift + row0_t + r;  // largest key index this row may see
      float m_ref = -INFINITY, l_run = 0.f;

      for (int j = 0; j < n_t; ++j, ++g) {
        const uint32_t buf = g & 1;
        const uint32_t s_col = s_col0 + buf * kBN;
        mbar_wait(bar(BAR_S_FULL + t * 2 + buf), (g >> 1) & 1);
        tc_fence_after();
        float s[kBN];
        {
          uint32_t raw[2][32];
          tmem_ld32(s_col, raw[0]);
          tmem_ld32(s_col + 32, raw[1]);
          tmem_wait_ld();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            s[i] = __uint_as_float(raw[0][i]);
            s[32 + i] = __uint_as_float(raw[1][i]);
          }
        }
        const int col0 = j * kBN;
        if (col0 + kBN - 1 > it.shift + row0_t) {  // step touches the diagonal / the end (uniform test)
          const int lim = q_limit - col0;
#pragma unroll
          for (int i = 0; i < kBN; ++i) s[i] = i <= lim ? s[i] : -INFINITY;
        }
        float mx4[4];  // four independent chains of 3-input max
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          mx4[q4] = fmaxf(s[16 * q4], s[16 * q4 + 1]);
#pragma unroll
          for (int i = 2; i < 16; i += 2) mx4[q4] = fmaxf(mx4[q4], fmaxf(s[16 * q4 + i], s[16 * q4 + i + 1]));
        }
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;
        // lazy rescale: keep exponentiating against m_ref until the true max outruns it by 2^8
        const bool grow = mx > m_ref + kRescaleThreshold;
        if (__any_sync(0xffffffffu, grow)) {
          const float m_new = grow ? mx : m_ref;
          if (j > 0) {
            const float sc = ex2_mufu(m_ref - m_new);  // 1 for lanes that keep their reference
            l_run *= sc;
            mbar_wait(bar(BAR_O_DONE + t), (g - 1) & 1);  // P V of the previous step retired: O is quiescent
            tc_fence_after();
#pragma unroll
            for (int c = 0; c < kD; c += 32) {
              uint32_t o[32];
              tmem_ld32(o_col + c, o);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * sc);
              tmem_st32(o_col + c, o);
            }
          }
          m_ref = m_new;
        }
        if (j == n_t - 1 && quarter == 0) {
          // keys >= Lk in the last block were never written: their V rows may hold NaN/Inf bit patterns and
          // 0 * NaN would poison O.  Zero those inside the last 16-key group the MMA reads (whole groups
          // past it are skipped by the issuer).
          const int stop = min(kBN, (kv_end_t - col0 + 15) & ~15);
          const int start = it.lk - col0;
          if (start < stop) {
            const uint32_t cnt = kv_base + j;
            const int st = cnt % kStages;
            mbar_wait(bar(BAR_V_FULL + st), (cnt / kStages) & 1);
            uint8_t* vg = smem_gen + (sV - smem_base) + st * kKVTileBytes;
            for (int row = start; row < stop; ++row)
              if (lane < 8 * kSlabs)
                *reinterpret_cast<uint4*>(vg + (lane >> 3) * kKVHalfBytes + row * 128 + (lane & 7) * 16) =
                    make_uint4(0, 0, 0, 0);
            fence_proxy_async_smem();
          }
        }
        const float m_use = m_ref == -INFINITY ? 0.f : m_ref;
        const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);
        const uint64_t negm2 = pack2(-m_use, -m_use);
        uint64_t sum2[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};  // two chains: FADD2 latency off the critical path
#pragma unroll
        for (int chunk = 0; chunk < kBN / 32; ++chunk) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const int col = chunk * 32 + i;
            const uint64_t x = fma2(pack2(s[col], s[col + 1]), scale2, negm2);
            float p0, p1;
            // kEmu of every 64 columns take the FMA-pipe exponential, interleaved with the MUFU ones
            if (kEmu > 0 && ((col / 2) * (kEmu / 2)) % (kBN / 2) < kEmu / 2) {
              ex2_poly2(x, p0, p1);
            } else {
              float x0, x1;
              unpack2(x, x0, x1);
              p0 = ex2_mufu(x0);
              p1 = ex2_mufu(x1);
            }
            sum2[(i >> 1) & 1] = add2(sum2[(i >> 1) & 1], pack2(p0, p1));
            pk[i / 2] = Mma16816<T>::pack(p0, p1);
          }
          tmem_st16(s_col + chunk * 16, pk);  // bf16 P overwrites the first 32 columns of this S buffer
        }
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(BAR_P_FULL + t * 2 + buf));
        float sum_lo, sum_hi;
        unpack2(add2(sum2[0], sum2[1]), sum_lo, sum_hi);
        l_run += sum_lo + sum_hi;
      }

      // ---- epilogue: O / l -> global ----
      // (not BAR_O_DONE: with S two steps ahead this warpgroup may be two P V commits behind, and a parity wait
      // can only tell one phase from the next)
      mbar_wait(bar(BAR_O_FINAL + t), items_done & 1);
      ++items_done;
      tc_fence_after();
      // a row that sees no key (Lk < Lq) still sums 2^-126 per FMA-pipe exponential: it must produce zeros
      const float inv = (l_run > 0.f && q_limit >= 0) ? 1.f / l_run : 0.f;
#pragma unroll
      for (int c = 0; c < kD; c += 32) {
        uint32_t o[32];
        tmem_ld32(o_col + c, o);
        tmem_wait_ld();
        if (r < rows_t) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 v;
            v.x = Mma16816<T>::pack(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
            v.y = Mma16816<T>::pack(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
            v.z = Mma16816<T>::pack(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
            v.w = Mma16816<T>::pack(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + c + i) = v;
          }
        }
      }
      tc_fence_before();  // the next item's first P V (accumulate = 0) overwrites O after our next P_FULL arrival
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

template <typename T, int kEmu, int kD>
static int launch(const PrefillArgs& a, cudaStream_t stream) {
  constexpr int kQTileBytes = (kD / 64) * kQHalfBytes;
  constexpr int kKVTileBytes = (kD / 64) * kKVHalfBytes;
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
  p.num_seqs = a.num_seqs;
  p.pages_per_block = paged ? a.block_size / kBN : 1;
  p.max_blocks = paged ? a.max_blocks : 1;
  p.num_blocks = paged ? (int32_t)a.num_blocks : 0;
  p.paged = paged ? 1 : 0;
  p.levels = (a.max_seqlen_q + kQTiles * kBM - 1) / (kQTiles * kBM);
  p.max_seqlen_q = a.max_seqlen_q;
  p.scale_log2 = a.scale * kLog2e;

  constexpr int smem = kQTiles * kQTileBytes + 2 * kStages * kKVTileBytes + 8 * BAR_COUNT + 16 + 1024;
  auto kern = prefill_dbs_kernel<T, kEmu, kD>;
  static SmemOptIn optin;  // one per instantiation
  if (int rc = optin_dynamic_smem((const void*)kern, smem, &optin, "prefill_dbs")) return rc;
  const int num_sms = device_sm_count();
  const long long items = (long long)p.levels * a.num_seqs * a.num_heads;
  const int grid = (int)(items < num_sms ? items : num_sms);  // persistent: one CTA per SM
  kern<<<grid, kThreads, smem, stream>>>(mq, mk, mv, p);
  return check_launch("prefill_dbs_kernel");
}

}  // namespace dbs

// returns -1 when the request is outside what this kernel covers
int prefill_dbs_dispatch(const PrefillArgs& a, cudaStream_t stream) {
  if (a.dtype != MVB200_BF16 && a.dtype != MVB200_F16) return -1;
  if (a.head_dim != 128 && a.head_dim != 64) return -1;
  if (a.max_seqlen_q <= 0) return -1;
  const long long levels = (a.max_seqlen_q + 255) / 256;
  if (levels > dbs::kMaxLevels || levels * a.num_seqs * a.num_heads >= (1LL << 31)) return -1;
  auto al16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  if (!al16(a.q) || !al16(a.out) || (a.q_row_stride % 8) || (a.out_row_stride % 8)) return -1;
  if (a.block_tables) {
    if (a.block_size % dbs::kBN || !al16(a.k_cache) || !al16(a.v_cache) || a.num_blocks >= (1LL << 31))
      return -1;
  } else {
    if (!al16(a.k) || !al16(a.v) || (a.k_row_stride % 8) || (a.v_row_stride % 8) || a.num_kv_rows <= 0)
      return -1;
  }
  note_kernel("prefill_tcgen05_dbs");
  const bool bf = a.dtype == MVB200_BF16;
#ifdef MVB200_TUNING
  static const int emu = [] {
    const char* e = getenv("MVB200_PREFILL_EMU");
    const int v = e ? atoi(e) : dbs::kDefaultEmu;
    return (v == 0 || v == 8 || v == 16 || v == 24 || v == 32) ? v : dbs::kDefaultEmu;
  }();
#define MVB_DBS_CASE(E) \
  if (emu == E && a.head_dim == 128) return bf ? dbs::launch<__nv_bfloat16, E, 128>(a, stream) : dbs::launch<__half, E, 128>(a, stream);
  MVB_DBS_CASE(0) MVB_DBS_CASE(8) MVB_DBS_CASE(16) MVB_DBS_CASE(24) MVB_DBS_CASE(32)
#undef MVB_DBS_CASE
#endif
  if (a.head_dim == 64)  // GPT-2 / OPT / Qwen2-0.5B heads
    return bf ? dbs::launch<__nv_bfloat16, dbs::kDefaultEmu, 64>(a, stream) : dbs::launch<__half, dbs::kDefaultEmu, 64>(a, stream);
  return bf ? dbs::launch<__nv_bfloat16, dbs::kDefaultEmu, 128>(a, stream)
            : dbs::launch<__half, dbs::kDefaultEmu, 128>(a, stream);
}

}  // namespace mvb200
