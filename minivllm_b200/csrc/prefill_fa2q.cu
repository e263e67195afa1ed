// prefill_fa2q.cu — persistent varlen causal flash attention, two query tiles per CTA (sm_100a).
//
// Replaces flash_attn_varlen_func(..., block_table=) at reference
// minivllm/models/layers/attention.py:422-433 (and its eager twin :467-504).
//
// Why a second kernel: the pipeline trace of prefill_tcgen05.cu (profiles/r01_prefill_v1_trace.log)
// shows 4.3 us of fixed cost per CTA, a 13 % tail, and Q K^T with 64-key tiles running at 2/3 rate
// because an SS MMA with N=64 needs 192 B/clk of shared-memory operand reads (probes/umma_pacing).
// This kernel removes all three:
//   * one persistent CTA per SM walks a longest-first list of work items
//     (sequence, head, 256-row query block) — no per-tile prologue, K/V prefetch runs across items;
//   * 128-key tiles: S = Q K^T is M128 N128 (128 B/clk of operand reads = the smem limit, full rate);
//   * two 128-row query tiles share every K/V tile (half the L2->smem traffic per FLOP) and
//     ping-pong on the tensor pipe: while softmax warpgroup 0 works on S0 the pipe runs
//     P1 V and Q1 K^T, and vice versa.
// Roles: warp 0 TMA producer, warp 1 MMA issuer (both warp-uniform loops, one elected lane issues),
// warps 2-3 idle (they keep the softmax warps on warp % 4 = TMEM lane quarter), then kHalves softmax
// warpgroups per query tile (thread = query row x a 128/kHalves-column slice of the score tile).
//   kHalves = 1 (384 threads): warps 4-7 tile 0, warps 8-11 tile 1, 128 scores per thread.
//   kHalves = 2 (640 threads): warps 4-7 / 8-11 the two column halves of tile 0, 12-15 / 16-19 of tile 1.
//     The softmax of a tile sits on the critical path  S(j) -> softmax -> P(j) V -> S(j+1)  (P aliases S in
//     TMEM) and a lone warp per SM sub-partition issues at ~0.4 IPC (profiles/r01_prefill_v2a_trace_emu32.log:
//     2100 of a 3500-clock period); two warps per sub-partition on the same tile interleave their MUFU and
//     FMA work.  The halves exchange their row maxima through shared memory (named barrier of the two warps
//     that share 32 rows) and keep separate row sums, added in the epilogue.
// TMEM (512 columns): S0 [0,128) S1 [128,256) O0 [256,384) O1 [384,512); bf16 P overwrites S in place.
// Tensor-bound: 4*D flops per unmasked (q,k) pair per head.
#include <math.h>
#include <stdlib.h>

#include <type_traits>

#include "prefill_args.h"
#include "softmax_math.cuh"
#include "tcgen05.cuh"

namespace mvb200 {
namespace fa2q {

using pf::add2;
using pf::ex2_mufu;
using pf::ex2_poly2;
using pf::fma2;
using pf::pack2;
using pf::unpack2;

// Pipeline trace (debug builds only, -DPF_TRACE; tools/prefill_trace.py): CTA 0 records clock64 at the
// hand-offs of its MMA issuer (role 0) and of one warp of each softmax warpgroup (roles 1, 2) for the
// first item it runs.
#ifdef PF_TRACE
__device__ unsigned long long g_trace2[4 * 64 * 8];
#define PF2_T(role, j, ev)                                                                    \
  do {                                                                                        \
    if (traced && (j) < 64) {                                                                 \
      unsigned long long t_;                                                                  \
      asm volatile("mov.u64 %0, %%clock64;" : "=l"(t_)::"memory");                            \
      g_trace2[((role) * 64 + (j)) * 8 + (ev)] = t_;                                          \
    }                                                                                         \
  } while (0)
#else
#define PF2_T(role, j, ev) do {} while (0)
#endif

constexpr int kD = 128;
constexpr int kBM = 128;                 // rows per query tile
constexpr int kQTiles = 2;               // query tiles per work item
constexpr int kBN = 128;                 // keys per K/V tile
constexpr int kPage = 64;                // keys per TMA box (= kvcache_block_size granularity)
constexpr int kStages = 2;               // K ring depth = V ring depth
constexpr int kDefaultHalves = 1;        // softmax warpgroups per query tile (2 measured slower: profiles/r02_prefill_ab_run5.jsonl)
constexpr int kMaxPChunks = 4;            // P is handed to the MMA warp in kPC (1, 2 or 4) column chunks per tile
constexpr int kTileHalfBytes = 128 * 128;  // [128 rows][128 B] = one 64-column half of a tile
constexpr int kTileBytes = 2 * kTileHalfBytes;
constexpr int kBoxBytes = kPage * 128;     // one TMA box: 64 rows x 64 columns
constexpr int kTmemCols = 512;
constexpr int kColS = 0, kColO = 256;      // + t * 128
constexpr int kMaxLevels = 512;            // query blocks per sequence (256 rows each): 131072 tokens
constexpr int kDefaultPChunks = 1;
constexpr int kDefaultEmu = 32;            // of every 128 columns, exponentials done on the FMA pipe
constexpr float kRescaleThreshold = 8.f;   // log2 domain: rescale O only when the max grew by > 2^8

enum {
  BAR_Q_FULL = 0,    // [2]  TMA -> MMA warp t
  BAR_Q_EMPTY = 2,   // [2]  MMA warp t (commit after the item's last Q_t K^T) -> TMA
  BAR_K_FULL = 4,    // [kStages]
  BAR_K_EMPTY = 6,   // [kStages]  MMA (commit after the last Q_t K^T that reads the stage) -> TMA
  BAR_V_FULL = 8,
  BAR_V_EMPTY = 10,
  BAR_S_FULL = 12,   // [2]  MMA warp t -> softmax warpgroup t
  BAR_O_DONE = 14,   // [2]  MMA warp t (commit after every P_t V) -> softmax warpgroup t
  BAR_P_FULL = 16,   // [2][kMaxPChunks]  softmax warpgroup t (4 warp arrivals) -> MMA warp t, per column chunk of P
  BAR_COUNT = 16 + 2 * kMaxPChunks
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
  int n[kQTiles];      // 128-key tiles tile t needs (0: nothing to attend to)
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

MVB_DEV void named_bar_sync(int id, int threads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}
template <int kRegs>
MVB_DEV void setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegs)); }
template <int kRegs>
MVB_DEV void setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegs)); }

template <typename T, int kEmu, int kPC, int kHalves>
__global__ void __launch_bounds__(128 + 256 * kHalves, 1)
prefill_fa2q_kernel(const __grid_constant__ CUtensorMap tmap_q,
                    const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const Params p) {
  constexpr bool kBf16 = sizeof(T) == 2 && !std::is_same<T, __half>::value;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const uint32_t sQ = smem_base;                         // [2 tiles][32 KB]
  const uint32_t sK = sQ + kQTiles * kTileBytes;         // [kStages][32 KB]
  const uint32_t sV = sK + kStages * kTileBytes;
  const uint32_t bars = sV + kStages * kTileBytes;
  const uint32_t tmem_slot = bars + 8 * BAR_COUNT;
  // row maxima [parity][tile][half][row] and row sums [tile][half][row] the column halves exchange
  float* xmax = reinterpret_cast<float*>(smem_gen + (tmem_slot + 16 - smem_base));
  float* xsum = xmax + 2 * kQTiles * 2 * kBM;
  auto bar = [&](int i) { return bars + 8u * i; };
  static_assert(kHalves == 1 || (kHalves == 2 && kPC == 1), "column halves hand P over in one piece");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_items = p.levels * p.num_seqs * p.num_heads;

  if (warp == 0) {
    tmem_alloc(tmem_slot, kTmemCols);
    tmem_relinquish();
  }
  if (threadIdx.x == 32) {
    for (int i = 0; i < BAR_COUNT; ++i) {
      mbar_init(bar(i), i >= BAR_P_FULL ? 4 * kHalves : 1);
    }
    fence_mbar_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem_gen + (tmem_slot - smem_base));
  if (warp < 4) {
    // one setmaxnreg for the whole warpgroup 0 (TMA, MMA and the two idle warps): the CTA pool is its launch allocation, 640 x 96: 128 x 64 + 512 x 104 = 61440
    if constexpr (kHalves == 2) setmaxnreg_dec<64>();
  }
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
          mbar_arrive_expect_tx(bar(BAR_Q_FULL + t), kTileBytes);
          const uint32_t dst = sQ + t * kTileBytes;
          tma_load_3d(dst, &tmap_q, bar(BAR_Q_FULL + t), 0, it.head, it.q0 + it.row0[t]);
          tma_load_3d(dst + kTileHalfBytes, &tmap_q, bar(BAR_Q_FULL + t), 64, it.head, it.q0 + it.row0[t]);
        }
        __syncwarp();
      }
      const int32_t* bt = p.block_tables + (int64_t)it.seq * p.bt_row_stride;
      const int kv_need = max(it.kv_end[0], it.kv_end[1]);
      for (int j = 0; j < it.n_kt; ++j, ++kv_cnt) {
        const int s = kv_cnt % kStages;
        const uint32_t ph = (kv_cnt / kStages) & 1;
        // the two 64-key pages of this tile: block id (paged) or packed row (dense)
        int c_blk[2], c_sub[2];
        int npages = 0;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int pg = 2 * j + h;
          if (pg * kPage >= kv_need) break;  // nobody reads past the last needed key
          ++npages;
          if (p.paged) {
            const int v = __ldg(bt + min(pg / p.pages_per_block, p.max_blocks - 1));
            c_blk[h] = (v >= 0 && v < p.num_blocks) ? v : p.num_blocks;  // invalid -> OOB -> zero fill
            c_sub[h] = (pg % p.pages_per_block) * kPage;
          } else {
            c_blk[h] = it.k0 + pg * kPage;
            c_sub[h] = 0;
          }
        }
        mbar_wait(bar(BAR_K_EMPTY + s), ph ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(bar(BAR_K_FULL + s), npages * 2 * kBoxBytes);
          for (int h = 0; h < npages; ++h) {
            const uint32_t dst = sK + s * kTileBytes + h * kBoxBytes;
            if (p.paged) {
              tma_load_4d(dst, &tmap_k, bar(BAR_K_FULL + s), 0, it.kvh, c_sub[h], c_blk[h]);
              tma_load_4d(dst + kTileHalfBytes, &tmap_k, bar(BAR_K_FULL + s), 64, it.kvh, c_sub[h], c_blk[h]);
            } else {
              tma_load_3d(dst, &tmap_k, bar(BAR_K_FULL + s), 0, it.kvh, c_blk[h]);
              tma_load_3d(dst + kTileHalfBytes, &tmap_k, bar(BAR_K_FULL + s), 64, it.kvh, c_blk[h]);
            }
          }
        }
        __syncwarp();
        mbar_wait(bar(BAR_V_EMPTY + s), ph ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(bar(BAR_V_FULL + s), npages * 2 * kBoxBytes);
          for (int h = 0; h < npages; ++h) {
            const uint32_t dst = sV + s * kTileBytes + h * kBoxBytes;
            if (p.paged) {
              tma_load_4d(dst, &tmap_v, bar(BAR_V_FULL + s), 0, it.kvh, c_sub[h], c_blk[h]);
              tma_load_4d(dst + kTileHalfBytes, &tmap_v, bar(BAR_V_FULL + s), 64, it.kvh, c_sub[h], c_blk[h]);
            } else {
              tma_load_3d(dst, &tmap_v, bar(BAR_V_FULL + s), 0, it.kvh, c_blk[h]);
              tma_load_3d(dst + kTileHalfBytes, &tmap_v, bar(BAR_V_FULL + s), 64, it.kvh, c_blk[h]);
            }
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    // Each query tile is a serial chain  S_t(j) -> softmax -> P_t(j) V(j) -> S_t(j+1)  (S_t(j+1) overwrites
    // P_t(j) in TMEM); the two chains interleave on the one tensor pipe, block by block.  P arrives in kPC
    // column chunks, so P V starts while the warpgroup still exponentiates.  Measured alternatives, both
    // slower: one issuing warp per tile (their MMAs interleave one by one and both chains stretch) and
    // a polling event loop over both chains (a poll pass costs ~500 cycles next to the softmax warps).
    constexpr uint32_t idesc_qk = umma_idesc(kBM, kBN, kBf16, false, false);
    constexpr uint32_t idesc_pv = umma_idesc(kBM, kD, kBf16, false, true);
    constexpr uint32_t desc_hi = umma_desc_hi(1024);  // 8-row (K-major) / 8-key (MN-major) groups 1 KiB apart
    const uint32_t q_lo0 = umma_desc_lo(sQ, 16);
    const uint32_t k_lo0 = umma_desc_lo(sK, 16);
    const uint32_t v_lo0 = umma_desc_lo(sV, kTileHalfBytes);  // LBO = distance between the 64-column halves
    uint32_t q_uses[kQTiles] = {0, 0};
    uint32_t p_cnt[kQTiles] = {0, 0};   // P tiles consumed so far (phase of the P_FULL barriers)
    uint32_t kv_cnt = 0;

    // S_t = Q_t K_s^T : 8 k-steps of 16 elements (32 B inside the 128-B swizzle row, halves 16 KiB apart)
    auto issue_qk = [&](int t, int s) {
      const uint32_t q_lo = q_lo0 + t * (kTileBytes >> 4);
      const uint32_t k_lo = k_lo0 + s * (kTileBytes >> 4);
      const uint32_t s_col = tmem + kColS + t * kBN;
#pragma unroll
      for (int kk = 0; kk < kD / 16; ++kk) {
        const uint32_t off = ((kk >> 2) * kTileHalfBytes + (kk & 3) * 32) >> 4;
        umma_ss(s_col, umma_desc(desc_hi, q_lo + off), umma_desc(desc_hi, k_lo + off), idesc_qk, kk > 0 ? 1u : 0u);
      }
    };

    for (int w = blockIdx.x; w < total_items; w += gridDim.x) {
      const Item it = decode_item(p, w);
      if (!it.valid || it.n_kt == 0) continue;
#ifdef PF_TRACE
      const bool traced = blockIdx.x == 0 && w == (int)blockIdx.x + (int)gridDim.x && lane == 0;
#endif
      PF2_T(0, 62, 0);
      // ---- prologue: S_t(0) ----
      {
        const int s = kv_cnt % kStages;
#pragma unroll
        for (int t = 0; t < kQTiles; ++t) {
          if (it.n[t] == 0) continue;
          mbar_wait(bar(BAR_Q_FULL + t), q_uses[t] & 1);
          ++q_uses[t];
        }
        mbar_wait(bar(BAR_K_FULL + s), (kv_cnt / kStages) & 1);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int t = 0; t < kQTiles; ++t) {
            if (it.n[t] == 0) continue;
            issue_qk(t, s);
            umma_commit(bar(BAR_S_FULL + t));
            if (it.n[t] == 1) umma_commit(bar(BAR_Q_EMPTY + t));
          }
          umma_commit(bar(BAR_K_EMPTY + s));
        }
        __syncwarp();
      }
      // ---- main loop: per K/V tile j and query tile t:  O_t += P_t(j) V(j) chunk by chunk, then
      //      S_t(j+1) = Q_t K(j+1)^T.  Blocking waits in chain order: the two chains run half a period apart,
      //      so while this warp follows tile t's chunks the other warpgroup is still exponentiating. ----
      for (int j = 0; j < it.n_kt; ++j) {
        const uint32_t cnt = kv_cnt + j;
        const int s = cnt % kStages, s_next = (cnt + 1) % kStages;
        mbar_wait(bar(BAR_V_FULL + s), (cnt / kStages) & 1);
        bool k_next_ready = false;
#pragma unroll
        for (int t = 0; t < kQTiles; ++t) {
          if (j >= it.n[t]) continue;
          const bool more = j + 1 < it.n[t];
          // last consumer of V(j) / K(j+1) among the query tiles: the highest t that still uses it
          const bool last_v = (t == kQTiles - 1) || (j >= it.n[t + 1 < kQTiles ? t + 1 : t]);
          const bool last_k = (t == kQTiles - 1) || (j + 1 >= it.n[t + 1 < kQTiles ? t + 1 : t]);
          const int ksteps = more ? kBN / 16 : (it.kv_end[t] - j * kBN + 15) / 16;
          const uint32_t v_lo = v_lo0 + s * (kTileBytes >> 4);
          const uint32_t p_col = tmem + kColS + t * kBN;
          const uint32_t o_col = tmem + kColO + t * kD;
#pragma unroll
          for (int ch = 0; ch < kPC; ++ch) {
            mbar_wait(bar(BAR_P_FULL + t * kMaxPChunks + ch), p_cnt[t] & 1);
            if (ch == 0) PF2_T(t, j, 1);
            tc_fence_after();
            if (elect_one()) {
              // A = bf16 P in TMEM (8 columns per 16 keys), B = V MN-major, 16 keys = 2 KiB
#pragma unroll
              for (int kk = ch * (8 / kPC); kk < (ch + 1) * (8 / kPC); ++kk) {
                if (kk < ksteps)
                  umma_ts(o_col, p_col + kk * 8, umma_desc(desc_hi, v_lo + kk * (16 * 128 >> 4)), idesc_pv,
                          (j > 0 || kk > 0) ? 1u : 0u);
              }
              if (ch == kPC - 1) {
                umma_commit(bar(BAR_O_DONE + t));
                if (last_v) umma_commit(bar(BAR_V_EMPTY + s));
              }
            }
            __syncwarp();
          }
          ++p_cnt[t];
          PF2_T(t, j, 4);
          if (more) {
            if (!k_next_ready) {
              mbar_wait(bar(BAR_K_FULL + s_next), ((cnt + 1) / kStages) & 1);
              k_next_ready = true;
              tc_fence_after();
            }
            if (elect_one()) {
              issue_qk(t, s_next);
              umma_commit(bar(BAR_S_FULL + t));
              if (j + 2 == it.n[t]) umma_commit(bar(BAR_Q_EMPTY + t));
              if (last_k) umma_commit(bar(BAR_K_EMPTY + s_next));
            }
            __syncwarp();
          }
          PF2_T(t, j, 6);
        }
      }
      kv_cnt += it.n_kt;
    }
  } else if (warp < 4) {
    // idle warps: keep the softmax warpgroups aligned to TMEM lane quarters (warp % 4)
  } else {
    // =========================== softmax warpgroups (thread = query row x column slice) ===========================
    // 640 threads start with 96 registers each (61440 of 65536): the softmax warpgroups take the rest
    if constexpr (kHalves == 2) setmaxnreg_inc<104>();
    constexpr int kCols = kBN / kHalves;   // score columns of a row this thread owns
    constexpr int kOCols = kD / kHalves;   // O columns this thread rescales / writes out
    const int sw = warp - 4;
    const int t = sw / (4 * kHalves);             // query tile of this warpgroup
    const int half = (sw >> 2) % kHalves;         // column slice
    const int quarter = warp & 3;                 // TMEM lanes this warp may touch
    const int r = quarter * 32 + lane;            // query row within the tile
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16);
    const uint32_t s_col = lane_base + kColS + t * kBN + half * kCols;
    const uint32_t p_col = lane_base + kColS + t * kBN + half * (kCols / 2);  // bf16 pairs: half the columns
    const uint32_t o_col = lane_base + kColO + t * kD + half * kOCols;
    const int pair_bar = 1 + t * 4 + quarter;     // named barrier of the two warps that own the same 32 rows
    T* out = reinterpret_cast<T*>(p.out);
    uint32_t c = 0;        // S tiles consumed by this warpgroup so far (= P V commits it may wait for)
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
      T* orow = out + (int64_t)(it.q0 + row0_t + r) * p.out_row_stride + (int64_t)it.head * kD + half * kOCols;
      if (n_t == 0) {  // Lk < Lq corner: rows that see no key produce zeros
        if (r < rows_t)
#pragma unroll
          for (int i = 0; i < kOCols; i += 8) *reinterpret_cast<uint4*>(orow + i) = make_uint4(0, 0, 0, 0);
        continue;
      }
#ifdef PF_TRACE
      const bool traced = blockIdx.x == 0 && w == (int)blockIdx.x + (int)gridDim.x && lane == 0 && quarter == 0 && half == 0;
#endif
      const int q_limit = it.shift + row0_t + r;  // largest key index this row may see
      float m_ref = -INFINITY, l_run = 0.f;       // l_run: this thread's columns only

      for (int j = 0; j < n_t; ++j, ++c) {
        PF2_T(2 + t, j, 0);
        mbar_wait(bar(BAR_S_FULL + t), c & 1);
        PF2_T(2 + t, j, 1);
        tc_fence_after();
        float s[kCols];
        {
          uint32_t raw[kCols / 32][32];
#pragma unroll
          for (int g = 0; g < kCols / 32; ++g) tmem_ld32(s_col + 32 * g, raw[g]);
          tmem_wait_ld();
          PF2_T(2 + t, j, 2);
#pragma unroll
          for (int g = 0; g < kCols / 32; ++g)
#pragma unroll
            for (int i = 0; i < 32; ++i) s[32 * g + i] = __uint_as_float(raw[g][i]);
        }
        const int col0 = j * kBN;
        if (col0 + kBN - 1 > it.shift + row0_t) {  // tile touches the diagonal / the end (uniform test)
          const int lim = q_limit - col0 - half * kCols;
#pragma unroll
          for (int i = 0; i < kCols; ++i) s[i] = i <= lim ? s[i] : -INFINITY;
        }
        float mx4[4];  // four independent chains of 3-input max (a single chain is 64 dependent ops)
#pragma unroll
        for (int q4 = 0; q4 < 4; ++q4) {
          constexpr int kQ = kCols / 4;
          mx4[q4] = fmaxf(s[kQ * q4], s[kQ * q4 + 1]);
#pragma unroll
          for (int i = 2; i < kQ; i += 2) mx4[q4] = fmaxf(mx4[q4], fmaxf(s[kQ * q4 + i], s[kQ * q4 + i + 1]));
        }
        float mxl = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        if constexpr (kHalves == 2) {
          // row maximum over both column halves.  Slots alternate with the tile parity: the partner reads
          // tile j's slot after barrier j and cannot be behind barrier j+1 when tile j+2 rewrites it.
          float* slot = xmax + ((c & 1) * kQTiles + t) * 2 * kBM;
          slot[half * kBM + r] = mxl;
          named_bar_sync(pair_bar, 64);  // also orders the partner's tcgen05.ld of S before our P stores below
          mxl = fmaxf(mxl, slot[(half ^ 1) * kBM + r]);
        }
        const float mx = mxl * p.scale_log2;
        // lazy rescale: keep exponentiating against m_ref until the true max outruns it by 2^8
        const bool grow = mx > m_ref + kRescaleThreshold;
        if (__any_sync(0xffffffffu, grow)) {
          const float m_new = grow ? mx : m_ref;
          if (j > 0) {
            const float sc = ex2_mufu(m_ref - m_new);  // 1 for lanes that keep their reference
            l_run *= sc;
            mbar_wait(bar(BAR_O_DONE + t), (c - 1) & 1);  // P V of the previous tile retired: O is quiescent
            tc_fence_after();
#pragma unroll
            for (int g = 0; g < kOCols; g += 16) {  // 16-column chunks: s[] is live, registers are tight
              uint32_t o[16];
              tmem_ld16(o_col + g, o);
              tmem_wait_ld();
#pragma unroll
              for (int i = 0; i < 16; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * sc);
              tmem_st16(o_col + g, o);
            }
          }
          m_ref = m_new;
        }
        if (j == n_t - 1 && quarter == 0 && half == 0) {
          // keys >= Lk in the last block were never written: their V rows may hold NaN/Inf bit patterns and
          // 0 * NaN would poison O.  Zero those inside the last 16-key group the MMA reads (whole groups
          // past it are skipped by the issuer).
          const int stop = min(kBN, (kv_end_t - col0 + 15) & ~15);
          const int start = it.lk - col0;
          if (start < stop) {
            const uint32_t cnt = kv_base + j;
            const int st = cnt % kStages;
            mbar_wait(bar(BAR_V_FULL + st), (cnt / kStages) & 1);
            uint8_t* vg = smem_gen + (sV - smem_base) + st * kTileBytes;
            for (int row = start; row < stop; ++row)
              if (lane < 16)
                *reinterpret_cast<uint4*>(vg + (lane >> 3) * kTileHalfBytes + row * 128 + (lane & 7) * 16) =
                    make_uint4(0, 0, 0, 0);
            fence_proxy_async_smem();
          }
        }
        const float m_use = m_ref == -INFINITY ? 0.f : m_ref;
        PF2_T(2 + t, j, 3);
        const uint64_t scale2 = pack2(p.scale_log2, p.scale_log2);
        const uint64_t negm2 = pack2(-m_use, -m_use);
        uint64_t sum2[2] = {pack2(0.f, 0.f), pack2(0.f, 0.f)};  // two chains: FADD2 latency off the critical path
#pragma unroll
        for (int chunk = 0; chunk < kCols / 32; ++chunk) {
          uint32_t pk[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const int col = chunk * 32 + i;
            const uint64_t x = fma2(pack2(s[col], s[col + 1]), scale2, negm2);
            float p0, p1;
            // kEmu of every 128 columns take the FMA-pipe exponential, interleaved with the MUFU ones
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
          tmem_st16(p_col + chunk * 16, pk);  // bf16 P overwrites the first 64 columns of S_t
          constexpr int kChunksPerPiece = (kCols / 32) / kPC;
          if ((chunk + 1) % kChunksPerPiece == 0) {  // hand this part of P to the MMA warp
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar(BAR_P_FULL + t * kMaxPChunks + chunk / kChunksPerPiece));
          }
        }
        float sum_lo, sum_hi;
        unpack2(add2(sum2[0], sum2[1]), sum_lo, sum_hi);
        l_run += sum_lo + sum_hi;
        PF2_T(2 + t, j, 4);

        PF2_T(2 + t, j, 6);
      }

      // ---- epilogue: O / l -> global ----
      float l_tot = l_run;
      if constexpr (kHalves == 2) {
        // (the next write of this slot is an item later, behind at least one pair barrier of that item)
        xsum[(t * 2 + half) * kBM + r] = l_run;
        named_bar_sync(pair_bar, 64);
        l_tot += xsum[(t * 2 + (half ^ 1)) * kBM + r];
      }
      mbar_wait(bar(BAR_O_DONE + t), (c - 1) & 1);
      tc_fence_after();
      // a row that sees no key (Lk < Lq) still sums 2^-126 per FMA-pipe exponential: it must produce zeros
      const float inv = (l_tot > 0.f && q_limit >= 0) ? 1.f / l_tot : 0.f;
#pragma unroll
      for (int g = 0; g < kOCols; g += 32) {
        uint32_t o[32];
        tmem_ld32(o_col + g, o);
        tmem_wait_ld();
        if (r < rows_t) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
            uint4 v;
            v.x = Mma16816<T>::pack(__uint_as_float(o[i + 0]) * inv, __uint_as_float(o[i + 1]) * inv);
            v.y = Mma16816<T>::pack(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
            v.z = Mma16816<T>::pack(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
            v.w = Mma16816<T>::pack(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
            *reinterpret_cast<uint4*>(orow + g + i) = v;
          }
        }
      }
      PF2_T(2 + t, 63, 0);
      tc_fence_before();  // the next item's first P V (accumulate = 0) overwrites O after our next P_FULL arrival
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, kTmemCols);
}

template <typename T, int kEmu, int kPC, int kHalves>
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
    const uint32_t box[4] = {64, 1, (uint32_t)kPage, 1};
    int rc = make_tensor_map(&mk, a.k_cache, 2, 4, dims, str, box, true);
    if (rc) return rc;
    rc = make_tensor_map(&mv, a.v_cache, 2, 4, dims, str, box, true);
    if (rc) return rc;
  } else {
    const uint64_t dims[3] = {(uint64_t)kD, (uint64_t)a.num_kv_heads, (uint64_t)a.num_kv_rows};
    const uint32_t box[3] = {64, 1, (uint32_t)kPage};
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
  p.pages_per_block = paged ? a.block_size / kPage : 1;
  p.max_blocks = paged ? a.max_blocks : 1;
  p.num_blocks = paged ? (int32_t)a.num_blocks : 0;
  p.paged = paged ? 1 : 0;
  p.levels = (a.max_seqlen_q + kQTiles * kBM - 1) / (kQTiles * kBM);
  p.max_seqlen_q = a.max_seqlen_q;
  p.scale_log2 = a.scale * kLog2e;

  constexpr int kThreads = 128 + 256 * kHalves;
  constexpr int smem = (kQTiles + 2 * kStages) * kTileBytes + 8 * BAR_COUNT + 16 + 3 * kQTiles * 2 * kBM * 4 + 1024;
  auto kern = prefill_fa2q_kernel<T, kEmu, kPC, kHalves>;
  static SmemOptIn optin;
  if (int rc = optin_dynamic_smem((const void*)kern, smem, &optin, "prefill_fa2q")) return rc;
  const int num_sms = device_sm_count();
  const long long items = (long long)p.levels * a.num_seqs * a.num_heads;
  const int grid = (int)(items < num_sms ? items : num_sms);  // persistent: one CTA per SM
  kern<<<grid, kThreads, smem, stream>>>(mq, mk, mv, p);
  return check_launch("prefill_fa2q_kernel");
}

}  // namespace fa2q

// returns -1 when the request is outside what this kernel covers
int prefill_fa2q_dispatch(const PrefillArgs& a, cudaStream_t stream) {
  if (a.dtype != MVB200_BF16 && a.dtype != MVB200_F16) return -1;
  if (a.head_dim != fa2q::kD) return -1;
  if (a.max_seqlen_q <= 0) return -1;
  const long long levels = (a.max_seqlen_q + 255) / 256;
  if (levels > fa2q::kMaxLevels || levels * a.num_seqs * a.num_heads >= (1LL << 31)) return -1;
  auto al16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  if (!al16(a.q) || !al16(a.out) || (a.q_row_stride % 8) || (a.out_row_stride % 8)) return -1;
  if (a.block_tables) {
    if (a.block_size % fa2q::kPage || !al16(a.k_cache) || !al16(a.v_cache) || a.num_blocks >= (1LL << 31))
      return -1;
  } else {
    if (!al16(a.k) || !al16(a.v) || (a.k_row_stride % 8) || (a.v_row_stride % 8) || a.num_kv_rows <= 0)
      return -1;
  }
  note_kernel("prefill_tcgen05_2q");
  const bool bf = a.dtype == MVB200_BF16;
#ifdef MVB200_TUNING
  // Tuning build only (python minivllm_b200/build.py --trace): columns of every 128 whose exponential
  // runs on the FMA pipe instead of MUFU (MVB200_PREFILL_EMU = 0 | 16 | 32 | 48) and the P hand-off
  // granularity (MVB200_PREFILL_PCHUNKS = 1 | 2 | 4).  The product library carries the measured optimum only.
  static const int emu = [] {
    const char* e = getenv("MVB200_PREFILL_EMU");
    const int v = e ? atoi(e) : fa2q::kDefaultEmu;
    return (v == 0 || v == 16 || v == 32 || v == 48 || v == 64) ? v : fa2q::kDefaultEmu;
  }();
  static const int pchunks = [] {
    const char* e = getenv("MVB200_PREFILL_PCHUNKS");
    const int v = e ? atoi(e) : fa2q::kDefaultPChunks;
    return (v == 1 || v == 2 || v == 4) ? v : fa2q::kDefaultPChunks;
  }();
  static const int halves = [] {
    const char* e = getenv("MVB200_PREFILL_HALVES");
    const int v = e ? atoi(e) : fa2q::kDefaultHalves;
    return (v == 1 || v == 2) ? v : fa2q::kDefaultHalves;
  }();
#define MVB_FA2Q_CASE(E, C)                                                                                   \
  if (emu == E && pchunks == C && halves == 1)                                                                \
    return bf ? fa2q::launch<__nv_bfloat16, E, C, 1>(a, stream) : fa2q::launch<__half, E, C, 1>(a, stream);
#define MVB_FA2Q_CASE2(E)                                                                                     \
  if (emu == E && halves == 2)                                                                                \
    return bf ? fa2q::launch<__nv_bfloat16, E, 1, 2>(a, stream) : fa2q::launch<__half, E, 1, 2>(a, stream);
  MVB_FA2Q_CASE(0, 1) MVB_FA2Q_CASE(0, 2) MVB_FA2Q_CASE(0, 4)
  MVB_FA2Q_CASE(16, 1) MVB_FA2Q_CASE(16, 2) MVB_FA2Q_CASE(16, 4)
  MVB_FA2Q_CASE(32, 1) MVB_FA2Q_CASE(32, 2) MVB_FA2Q_CASE(32, 4)
  MVB_FA2Q_CASE(48, 1) MVB_FA2Q_CASE(48, 2) MVB_FA2Q_CASE(48, 4)
  MVB_FA2Q_CASE2(0) MVB_FA2Q_CASE2(16) MVB_FA2Q_CASE2(32) MVB_FA2Q_CASE2(48) MVB_FA2Q_CASE2(64)
#undef MVB_FA2Q_CASE
#undef MVB_FA2Q_CASE2
  return fail(MVB200_EINVAL, "prefill: no kernel instance for emu %d / pchunks %d / halves %d", emu, pchunks, halves);
#else
  return bf ? fa2q::launch<__nv_bfloat16, fa2q::kDefaultEmu, fa2q::kDefaultPChunks, fa2q::kDefaultHalves>(a, stream)
            : fa2q::launch<__half, fa2q::kDefaultEmu, fa2q::kDefaultPChunks, fa2q::kDefaultHalves>(a, stream);
#endif
}

}  // namespace mvb200

#ifdef PF_TRACE
extern "C" __attribute__((visibility("default"))) int mvb200_debug_read_prefill2_trace(unsigned long long* dst) {
  return (int)cudaMemcpyFromSymbol(dst, mvb200::fa2q::g_trace2, sizeof(mvb200::fa2q::g_trace2));
}
#endif
