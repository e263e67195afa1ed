// paged_decode.cu — split-KV paged GQA decode attention for sm_100a.
//
// Replaces flash_attn_with_kvcache at reference minivllm/models/layers/attention.py:439-447 and
// the eager PageAttention.__call__ (minivllm/models/layers/page_attention.py:35-133).
//
// Fast path (16-bit K/V, D in {64,128}, block_size % 64 == 0, group <= 16):
//   Work unit = (sequence, KV split, kv head).  The grid is PERSISTENT: two resident CTAs per SM draw
//   units from a device-side counter (kv head fastest, so co-resident CTAs read adjacent 256-byte
//   segments of the same token rows) until none is left.  Per CTA a producer warp turns block-table
//   entries into TMA coordinates and issues cp.async.bulk.tensor loads of whole K and V tiles (64 tokens
//   x D, 128B-swizzled) into a multi-stage shared-memory ring guarded by full/empty mbarriers; the ring
//   keeps streaming across unit boundaries (the unit's Q rows ride along as a 1-D bulk copy).  Four
//   consumer warps each own 16 tokens of a tile: Q(group padded to 16 rows) x K^T and P x V run on
//   mma.sync m16n8k16 with ldmatrix-fed fragments (K/V are read from HBM once per kv head, not once per
//   query head), online softmax in fp32 with quad shuffles, exp2 with the scale folded in.  At the end
//   of a unit the warps merge through a ring slot the producer lends them.  With num_splits > 1 each
//   unit writes a normalised partial + log-sum-exp and counts itself on the ticket of its (sequence, kv
//   head); once the units run out the same CTAs draw (sequence, head) rows from a second counter and
//   combine each row's partials as soon as its ticket is complete (no second kernel, no launch gap).
//   HBM-bound: algorithmic bytes = ctx * D * 2 (K,V) * elem per (sequence, kv head).
//
// Generic path (fp32 caches, any D <= 256, any block size, any group): one CTA per (b, h),
//   CUDA-core dot products.  Used for the reference's CPU-sized unit-test shapes.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace mvb200 {

// =============================================================================================
// fast path: persistent CTAs over (sequence, split, kv head) units, fused split reduction
// =============================================================================================

constexpr int kTileTokens = 64;
constexpr int kConsumerWarps = 4;
constexpr int kConsumerThreads = kConsumerWarps * 32;
constexpr int kDecodeThreads = (kConsumerWarps + 1) * 32;
constexpr int kMaxGroup = 16;
constexpr int kQSlotBytes = kMaxGroup * 128 * 2;  // Q of one unit: up to 16 rows x 128 x 16 bit

// Persistent part of the workspace (the LAST kDecodeTailBytes of the caller's buffer, at the same place
// for every batch size / split count that shares the buffer): one arrival ticket per (sequence, kv head)
// and the dynamic work counters.  Zeroed once by mvb200_decode_workspace_init; the kernel leaves it zeroed.
constexpr int kDecodeMaxTickets = 65536;
constexpr int kSchedInts = 16;  // [0] next unit, [1] CTAs past their last unit, [3] cookie, [4]/[5] same for reduce rows
constexpr uint32_t kWorkspaceCookie = 0x6d766232u;
constexpr size_t kDecodeTailBytes = (size_t)kDecodeMaxTickets * 4 + kSchedInts * 4;

struct DecodeParams {
  const void* q;
  int64_t q_row_stride;  // elements between tokens
  const int32_t* block_tables;
  int64_t bt_row_stride;
  const int32_t* context_lens;
  void* out;
  int64_t out_row_stride;
  float* part_o;     // [B, H, S, D]
  float* part_lse;   // [B, H, S]  (log2 domain)
  int32_t* tickets;  // [B * KVH] arrivals of a (sequence, kv head)'s splits; nullptr when num_splits == 1
  int32_t* sched;    // dynamic unit counter (nullptr: static round-robin over the grid)
  int32_t batch;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t group;
  int32_t max_blocks;       // block-table width (MB)
  int32_t tiles_per_block;  // block_size / 64
  int32_t num_blocks;       // NB (also the out-of-bounds TMA coordinate -> zero fill)
  int32_t num_splits;
  int32_t stages;
  int32_t total_units;  // B * num_splits * KVH
  int32_t q_bulk;       // Q rows of a unit can be fetched with one 16-byte-aligned bulk copy
  int32_t early;        // MVB200_DECODE_KV_PREFIX_STABLE: K/V tiles may be fetched before the dependency wait
  // fused store of the step's new tokens (nullptr: the caller stored them): row b goes to slot_mapping[b]
  const void* new_k;
  const void* new_v;
  int64_t new_k_stride, new_v_stride;  // elements between token rows
  const int32_t* slot_mapping;
  void* k_cache;
  void* v_cache;
  int64_t num_slots;
  float scale_log2;
};

// What the producer tells the consumer warps about a unit (written to shared memory before the `full`
// arrival of the unit's first ring slot).
struct __align__(16) UnitInfo {
  int32_t b, kvh, split, t0;
  int32_t n;  // tiles of the unit; -1 = no more units
  int32_t ctx;
  int32_t pad[2];
};

MVB_DEV void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(bar)
      : "memory");
}
MVB_DEV void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumerThreads) : "memory"); }
MVB_DEV float ld_cg_f32(const float* p) {
  float v;
  asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

template <typename T, int D>
__global__ void __launch_bounds__(kDecodeThreads)
paged_decode_kernel(const __grid_constant__ CUtensorMap tmap_k,
                    const __grid_constant__ CUtensorMap tmap_v, const DecodeParams p) {
  constexpr int kHalves = D / 64;                       // 128-byte column groups per row
  constexpr int kHalfBytes = kTileTokens * 128;         // one [64 tokens][128 B] swizzled slab
  constexpr int kTileBytes = kHalves * kHalfBytes;      // K (or V) tile
  constexpr int kStageBytes = 2 * kTileBytes;           // K + V
  constexpr int kKSteps = D / 16;
  static_assert(kConsumerWarps * kMaxGroup * D * 4 <= kStageBytes, "merge scratch must fit the ring slot it borrows");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem base is only guaranteed 16B aligned: round up to the swizzle atom
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_gen = smem_raw + (smem_base - smem_u32(smem_raw));
  const int stages = p.stages;
  // layout: ring[stages] | q[stages] | full[stages] empty[stages] | UnitInfo[stages] | mm[64] ml[64]
  const uint32_t q_base = smem_base + stages * kStageBytes;
  const uint32_t bar_base = q_base + stages * kQSlotBytes;
  UnitInfo* uinfo = reinterpret_cast<UnitInfo*>(smem_gen + (bar_base - smem_base) + 16 * stages);
  float* mm = reinterpret_cast<float*>(uinfo + stages);  // [4][16] running max of each warp's rows
  float* ml = mm + kConsumerWarps * kMaxGroup;           // [4][16] running sum

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int G = p.group;
  const int S = p.num_splits;
  const int KVH = p.num_kv_heads;

  // prologue that needs nothing from earlier kernels runs before the dependency wait (PDL)
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(bar_base + 8 * s, 1);                           // full: producer's (expect_tx) arrive
      mbar_init(bar_base + 8 * (stages + s), kConsumerWarps);   // empty: one arrive per warp
    }
    fence_mbar_init();
  }
  if (threadIdx.x == kConsumerWarps * 32) {
    tma_prefetch_desc(&tmap_k);
    tma_prefetch_desc(&tmap_v);
  }
  __syncthreads();
  // q, context_lens, block tables and the freshly stored K/V are visible after the dependency wait.  With
  // p.early the caller vouches that the kernel launched just before this one writes only the NEWEST token of
  // each sequence (and q): the producer warp then starts streaming the older tiles of its first unit while
  // that kernel is still draining, and waits just before the first load that could see its writes.
  const bool run_ahead = p.early && warp == kConsumerWarps;
  if (!run_ahead) {
    pdl_wait();
    pdl_launch_dependents();  // the next kernel may start its own prologue
  }

  // Workspace stamp, checked by a consumer thread: it is past its dependency wait, so an init kernel launched just
  // before this one on the stream (first decode after mvb200_decode_workspace_init) is complete and visible.  A
  // run-ahead producer could read the stamp too early and trap on a workspace that is being initialised.
  if (p.sched != nullptr && blockIdx.x == 0 && threadIdx.x == 0 &&
      reinterpret_cast<volatile uint32_t*>(p.sched)[3] != kWorkspaceCookie) {
    printf("mvb200: decode workspace was not initialised (mvb200_decode_workspace_init)\n");
    __trap();
  }

  T* out = reinterpret_cast<T*>(p.out);

  if (warp == kConsumerWarps) {
    // ------------------------------ producer warp ------------------------------
    // Walks this CTA's units back to back: the ring keeps streaming across unit boundaries, so the
    // consumers' merge / write-out of one unit overlaps the loads of the next (no per-CTA prologue and
    // no wave tail as with one CTA per unit).  Ring entries of a unit: its n tiles, then one slot that
    // carries no data and is lent to the consumers as merge scratch.
    int it = 0;
    bool waited = !run_ahead;  // false: nothing the preceding kernel writes may be touched yet
    // The first unit is the CTA's own index (no round trip to the counter before the first load); the next
    // one is drawn when a unit STARTS, so the atomic's latency hides behind that unit's tiles.  The counter
    // therefore counts draws beyond the first round.  (The preceding launch re-arms the counter as it ends:
    // a run-ahead producer draws only after its dependency wait.)
    int u = blockIdx.x;
    for (;;) {
      if (u >= p.total_units) {
        if (!waited) {
          pdl_wait();
          pdl_launch_dependents();
          waited = true;
        }
        if (lane == 0) {
          const int s = it % stages;
          mbar_wait(bar_base + 8 * (stages + s), ((it / stages) & 1) ^ 1);
          uinfo[s].n = -1;
          mbar_arrive(bar_base + 8 * s);
          if (p.sched != nullptr) {
            // every CTA draws exactly one out-of-range unit: the last one to do so re-arms the counters
            const int done = atomicAdd(p.sched + 1, 1);
            if (done == (int)gridDim.x - 1) {
              p.sched[0] = 0;
              p.sched[1] = 0;
              __threadfence();
            }
          }
        }
        break;
      }
      int u_next = 0;
      bool drawn = false;
      if (waited) {
        if (lane == 0) u_next = p.sched != nullptr ? (int)gridDim.x + atomicAdd(p.sched, 1) : u + (int)gridDim.x;
        drawn = true;
      }
      const int kvh = u % KVH;
      const int rest = u / KVH;
      const int split = rest % S;
      const int b = rest / S;
      const int ctx = __ldg(p.context_lens + b);
      const int n_tiles = ctx > 0 ? (ctx + kTileTokens - 1) / kTileTokens : 0;
      const int tps = (n_tiles + S - 1) / S;
      const int t0 = split * tps;
      const int t1 = min(n_tiles, t0 + tps);
      const int n = max(t1 - t0, 0);
      const int32_t* bt = p.block_tables + (int64_t)b * p.bt_row_stride;
      const uint32_t q_bytes = (uint32_t)(G * D * (int)sizeof(T));
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
          const int idx = base + i;  // tile of the unit
          // Run-ahead ends before the sequence's last tile (it holds the token stored just now) and before
          // the ring would fill up: the consumers cannot free a slot until the unit's Q — which the
          // preceding kernels produce — has arrived on the first slot's barrier.
          const bool end_run_ahead = !waited && (t0 + idx == n_tiles - 1 || idx == stages || idx == n - 1);
          if (end_run_ahead && idx > 0) {
            pdl_wait();
            pdl_launch_dependents();
            waited = true;
            if (lane == 0 && p.q_bulk) {  // the Q copy the first slot's barrier is still expecting
              const int s0 = (it - idx) % stages;
              bulk_load_1d(q_base + s0 * kQSlotBytes,
                           reinterpret_cast<const T*>(p.q) + (int64_t)b * p.q_row_stride + (int64_t)kvh * G * D,
                           q_bytes, bar_base + 8 * s0);
            }
          } else if (end_run_ahead) {
            pdl_wait();
            pdl_launch_dependents();
            waited = true;
          }
          if (lane == 0) {
            const int s = it % stages;
            const uint32_t full = bar_base + 8 * s;
            mbar_wait(bar_base + 8 * (stages + s), ((it / stages) & 1) ^ 1);
            const bool first = idx == 0;
            if (first) uinfo[s] = UnitInfo{b, kvh, split, t0, n, ctx, {0, 0}};
            mbar_arrive_expect_tx(full, kStageBytes + (first && p.q_bulk ? q_bytes : 0u));
            if (first && p.q_bulk && waited)
              bulk_load_1d(q_base + s * kQSlotBytes,
                           reinterpret_cast<const T*>(p.q) + (int64_t)b * p.q_row_stride + (int64_t)kvh * G * D,
                           q_bytes, full);
            const uint32_t kdst = smem_base + s * kStageBytes;
            const uint32_t vdst = kdst + kTileBytes;
#pragma unroll
            for (int h = 0; h < kHalves; ++h) {
              tma_load_4d(kdst + h * kHalfBytes, &tmap_k, full, h * 64, kvh, sub, bid);
              tma_load_4d(vdst + h * kHalfBytes, &tmap_v, full, h * 64, kvh, sub, bid);
            }
          }
          ++it;
          __syncwarp();
        }
      }
      if (!waited) {  // an empty first unit
        pdl_wait();
        pdl_launch_dependents();
        waited = true;
      }
      if (!drawn && lane == 0)
        u_next = p.sched != nullptr ? (int)gridDim.x + atomicAdd(p.sched, 1) : u + (int)gridDim.x;
      // the scratch slot: a ring entry without data (4 warps x G rows x D floats = 16 G D bytes never
      // exceed a stage of 256 D bytes at G <= 16)
      if (lane == 0) {
        const int s = it % stages;
        mbar_wait(bar_base + 8 * (stages + s), ((it / stages) & 1) ^ 1);
        if (n == 0) uinfo[s] = UnitInfo{b, kvh, split, t0, 0, ctx, {0, 0}};
        mbar_arrive(bar_base + 8 * s);
      }
      ++it;
      __syncwarp();
      u = __shfl_sync(0xffffffffu, u_next, 0);
    }
    return;
  }

  // ------------------------------ consumer warps -----------------------------
  const int ctid = threadIdx.x;  // 0..127
  // ldmatrix lane geometry (see DESIGN.md "decode fragment map")
  const int krow = 16 * warp + (lane & 7) + ((lane >> 4) << 3);
  const int ksel = (lane >> 3) & 1;
  const int vrow = 16 * warp + (lane & 7) + (((lane >> 3) & 1) << 3);
  const int vsel = lane >> 4;
  int it = 0;
  for (;;) {
    {
      const int s = it % stages;
      mbar_wait(bar_base + 8 * s, (it / stages) & 1);
    }
    const UnitInfo info = uinfo[it % stages];
    if (info.n < 0) break;
    const int b = info.b, kvh = info.kvh, split = info.split, t0 = info.t0, n = info.n, ctx = info.ctx;

    // Fused store_kvcache (attention_backend.py:237-266): the unit that reads a sequence's last tile also owns
    // the step's new token of its kv head.  The warp whose 16 tokens include it fetches the new K/V row segment
    // at the start of the unit (one 16-byte vector per lane: K on the low lanes, V above), and when the last tile
    // has landed writes it INTO THE TILE in shared memory (the TMA copy of that row is stale) and to its slot in
    // the cache — plain stores, no fence on the load path: nobody else reads that row in this launch.
    constexpr int kVecsPerSeg = D * (int)sizeof(T) / 16;
    const int seq_tiles = ctx > 0 ? (ctx + kTileTokens - 1) / kTileTokens : 0;
    const int new_row = (ctx - 1) & (kTileTokens - 1);
    const bool own_new = p.new_k != nullptr && (n > 0 ? (t0 + n == seq_tiles && warp == (new_row >> 4))
                                                      : (seq_tiles == 0 && split == 0 && warp == 0));
    int64_t new_slot = -1;
    uint4 new_vec = make_uint4(0, 0, 0, 0);
    if (own_new) {
      new_slot = __ldg(p.slot_mapping + b);
      if (new_slot < 0 || new_slot >= p.num_slots) new_slot = -1;
      if (new_slot >= 0 && lane < 2 * kVecsPerSeg) {
        const bool is_v = lane >= kVecsPerSeg;
        const T* src = is_v ? reinterpret_cast<const T*>(p.new_v) + (int64_t)b * p.new_v_stride
                            : reinterpret_cast<const T*>(p.new_k) + (int64_t)b * p.new_k_stride;
        new_vec = __ldg(reinterpret_cast<const uint4*>(src + (int64_t)kvh * D) + (is_v ? lane - kVecsPerSeg : lane));
        if (n == 0) {  // a row without context: nothing to attend to, the cache still gets the token
          T* dst = reinterpret_cast<T*>(is_v ? p.v_cache : p.k_cache) + (new_slot * KVH + kvh) * D;
          reinterpret_cast<uint4*>(dst)[is_v ? lane - kVecsPerSeg : lane] = new_vec;
        }
      }
    }

    float o_acc[D / 8][4];
    float m_run[2] = {-INFINITY, -INFINITY};
    float l_run[2] = {0.f, 0.f};
#pragma unroll
    for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;

    if (n > 0) {
      // Q fragments (A operand, 16 x D, rows >= group are zero): from the unit's Q slot, or from global
      uint32_t qf[kKSteps][4];
      {
        const int g = lane >> 2, t = lane & 3;
        if (p.q_bulk) {
          const uint8_t* qs = smem_gen + (q_base - smem_base) + (it % stages) * kQSlotBytes;
#pragma unroll
          for (int kk = 0; kk < kKSteps; ++kk) {
            const int c = (kk * 16 + 2 * t) * 2;
            qf[kk][0] = g < G ? *reinterpret_cast<const uint32_t*>(qs + g * D * 2 + c) : 0u;
            qf[kk][1] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(qs + (g + 8) * D * 2 + c) : 0u;
            qf[kk][2] = g < G ? *reinterpret_cast<const uint32_t*>(qs + g * D * 2 + c + 16) : 0u;
            qf[kk][3] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(qs + (g + 8) * D * 2 + c + 16) : 0u;
          }
        } else {
          const T* qb = reinterpret_cast<const T*>(p.q) + (int64_t)b * p.q_row_stride + (int64_t)kvh * G * D;
#pragma unroll
          for (int kk = 0; kk < kKSteps; ++kk) {
            const int c = kk * 16 + 2 * t;
            qf[kk][0] = g < G ? *reinterpret_cast<const uint32_t*>(qb + g * D + c) : 0u;
            qf[kk][1] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(qb + (g + 8) * D + c) : 0u;
            qf[kk][2] = g < G ? *reinterpret_cast<const uint32_t*>(qb + g * D + c + 8) : 0u;
            qf[kk][3] = g + 8 < G ? *reinterpret_cast<const uint32_t*>(qb + (g + 8) * D + c + 8) : 0u;
          }
        }
      }

      for (int j = 0; j < n; ++j, ++it) {
        const int s = it % stages;
        const uint32_t full = bar_base + 8 * s;
        const uint32_t empty = bar_base + 8 * (stages + s);
        const uint32_t kbase = smem_base + s * kStageBytes;
        const uint32_t vbase = kbase + kTileBytes;
        const int tile_tok0 = (t0 + j) * kTileTokens;
        const bool partial = tile_tok0 + kTileTokens > ctx;

        if (j > 0) mbar_wait(full, (it / stages) & 1);

        if (own_new && j == n - 1 && new_slot >= 0) {
          if (lane < 2 * kVecsPerSeg) {
            const bool is_v = lane >= kVecsPerSeg;
            const int c = is_v ? lane - kVecsPerSeg : lane;
            uint8_t* tile = smem_gen + ((is_v ? vbase : kbase) - smem_base);
            *reinterpret_cast<uint4*>(tile + (c >> 3) * kHalfBytes + new_row * 128 + (((c & 7) ^ (new_row & 7)) << 4)) = new_vec;
            T* dst = reinterpret_cast<T*>(is_v ? p.v_cache : p.k_cache) + (new_slot * KVH + kvh) * D;
            reinterpret_cast<uint4*>(dst)[c] = new_vec;
          }
          fence_proxy_async_smem();  // generic writes before the next TMA refill of this slot
          __syncwarp();
        }

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
          for (int jj = 0; jj < 2; ++jj) {
            if (c + 8 * jj >= valid) sacc[jj][0] = sacc[jj][2] = -INFINITY;
            if (c + 8 * jj + 1 >= valid) sacc[jj][1] = sacc[jj][3] = -INFINITY;
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
          for (int jj = 0; jj < 2; ++jj) {
            sacc[jj][2 * r] = exp2f(sacc[jj][2 * r] * p.scale_log2 - m_use);
            sacc[jj][2 * r + 1] = exp2f(sacc[jj][2 * r + 1] * p.scale_log2 - m_use);
            sum += sacc[jj][2 * r] + sacc[jj][2 * r + 1];
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
      // the unit's scratch slot: a ring entry without data
      mbar_wait(bar_base + 8 * (it % stages), (it / stages) & 1);
    }

    // ------------------------------ merge the four warps ------------------------------
    // scratch[w][row][d ^ ((row & 7) << 3)] in the borrowed slot (XOR keeps the float2 stores of the eight
    // row groups of a warp off each other's banks)
    float* mo = reinterpret_cast<float*>(smem_gen + (size_t)(it % stages) * kStageBytes);
    {
      const int g = lane >> 2, t = lane & 3;
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int row = g + 8 * r;
        if (row < G) {
          float* dst = mo + (warp * G + row) * D;
#pragma unroll
          for (int i = 0; i < D / 8; ++i)
            *reinterpret_cast<float2*>(dst + ((i ^ (row & 7)) << 3) + 2 * t) = make_float2(o_acc[i][2 * r], o_acc[i][2 * r + 1]);
          if (t == 0) {
            mm[warp * kMaxGroup + row] = m_run[r];
            ml[warp * kMaxGroup + row] = l_run[r];
          }
        }
      }
    }
    consumer_sync();
    for (int idx = ctid; idx < G * D; idx += kConsumerThreads) {
      const int row = idx / D, d = idx - row * D;
      const int ds = (((d >> 3) ^ (row & 7)) << 3) + (d & 7);
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < kConsumerWarps; ++w) M = fmaxf(M, mm[w * kMaxGroup + row]);
      const float Ms = M == -INFINITY ? 0.f : M;
      float L = 0.f, o = 0.f;
#pragma unroll
      for (int w = 0; w < kConsumerWarps; ++w) {
        const float sc = exp2f(mm[w * kMaxGroup + row] - Ms);
        L += ml[w * kMaxGroup + row] * sc;
        o += mo[(w * G + row) * D + ds] * sc;
      }
      const float val = L > 0.f ? o / L : 0.f;
      const int h = kvh * G + row;
      if (S == 1) {
        out[(int64_t)b * p.out_row_stride + (int64_t)h * D + d] = from_f32<T>(val);
      } else {
        const int64_t ph = ((int64_t)b * p.num_heads + h) * S + split;
        p.part_o[ph * D + d] = val;
        if (d == 0) p.part_lse[ph] = L > 0.f ? M + log2f(L) : -INFINITY;
      }
    }

    if (S > 1) {
      // publish: bar.sync orders the CTA's partial stores before thread 0, whose fence (cumulative) makes
      // them visible device-wide before the ticket is counted
      if (p.tickets != nullptr) {
        consumer_sync();
        if (ctid == 0) {
          __threadfence();
          atomicAdd(p.tickets + ((int64_t)b * KVH + kvh), 1);
        }
      }
    }

    // hand the borrowed slot back
    consumer_sync();
    if (lane == 0) mbar_arrive(bar_base + 8 * (stages + (it % stages)));
    ++it;
  }

  if (S == 1 || p.tickets == nullptr) return;  // no tickets: decode_reduce_kernel follows
  // ------------------------------ phase 2: split reduction, same kernel ------------------------------
  // The consumer warps of every CTA now take (sequence, head) rows — the CTA's own index first, then from a
  // second counter — and combine a row's S partials as soon as all splits of its (sequence, kv head) have been
  // counted on the ticket: the units were handed out in the same order, so the wait is short, and whoever
  // still owes a partial is running, not queued behind this CTA (a CTA gets here only after every unit was
  // drawn).  The four warps take interleaved splits (partial rows of different splits in flight together: a
  // serial walk over 74 splits costs more than streaming the KV at batch 1), each against the maximum of its
  // own splits; lanes cover head_dim with 16-byte loads.  The idle ring is the merge scratch.
  // Tickets and counters are re-armed for the next launch by their last user.
  constexpr int kVecs = D / 4;  // float4 per partial row (32 for D = 128)
  const int total_rows = p.batch * p.num_heads;
  float4* racc = reinterpret_cast<float4*>(smem_gen);                // [4 warps][32 lanes]
  float* rml = reinterpret_cast<float*>(smem_gen + 4 * 32 * 16);     // [4] max, [4] weight sum
  int* rnext = reinterpret_cast<int*>(smem_gen + 4 * 32 * 16 + 64);  // next row of this CTA
  int r = blockIdx.x;
  for (;;) {
    if (r >= total_rows) {
      if (ctid == 0) {
        const int done = atomicAdd(p.sched + 5, 1);  // every CTA ends on exactly one out-of-range row
        if (done == (int)gridDim.x - 1) {
          p.sched[4] = 0;
          p.sched[5] = 0;
          __threadfence();
        }
      }
      break;
    }
    const int b = r / p.num_heads, h = r % p.num_heads;
    int32_t* ticket = p.tickets + b * KVH + h / G;
    if (ctid == 0) {
      *rnext = (int)gridDim.x + atomicAdd(p.sched + 4, 1);  // latency hides behind this row
      int seen;
      do {
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(ticket) : "memory");
        if (seen < S) __nanosleep(32);
      } while (seen < S);
    }
    consumer_sync();
    const int64_t ph = (int64_t)r * S;
    // this warp's splits: warp, warp + 4, ...  (S <= 128: at most 32 per warp, one per lane)
    const int mine = (S - warp + kConsumerWarps - 1) / kConsumerWarps;
    float wl = lane < mine ? __ldcg(p.part_lse + ph + warp + kConsumerWarps * lane) : -INFINITY;
    float M = wl;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, o));
    wl = wl == -INFINITY ? 0.f : exp2f(wl - M);  // an empty split has weight 0 (its row holds no data)
    const float W = warp_sum(wl);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4* rows = reinterpret_cast<const float4*>(p.part_o + (ph + warp) * D) + (lane < kVecs ? lane : 0);
#pragma unroll 8
    for (int j = 0; j < mine; ++j) {
      const float w = __shfl_sync(0xffffffffu, wl, j);
      float4 v = __ldcg(rows + (int64_t)(kConsumerWarps * j) * kVecs);
      if (w == 0.f) v = make_float4(0.f, 0.f, 0.f, 0.f);  // never-written row: whatever was read is dropped
      acc.x += w * v.x;
      acc.y += w * v.y;
      acc.z += w * v.z;
      acc.w += w * v.w;
    }
    racc[warp * 32 + lane] = acc;
    if (lane == 0) {
      rml[warp] = M;
      rml[4 + warp] = W;
    }
    consumer_sync();
    if (warp == 0) {
      const float Mx = fmaxf(fmaxf(rml[0], rml[1]), fmaxf(rml[2], rml[3]));
      float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
      float Wt = 0.f;
#pragma unroll
      for (int w4 = 0; w4 < kConsumerWarps; ++w4) {
        const float sc = rml[w4] == -INFINITY ? 0.f : exp2f(rml[w4] - Mx);
        const float4 a = racc[w4 * 32 + lane];
        o.x += sc * a.x;
        o.y += sc * a.y;
        o.z += sc * a.z;
        o.w += sc * a.w;
        Wt += sc * rml[4 + w4];
      }
      if (lane < kVecs) {
        const float inv = Wt > 0.f ? 1.f / Wt : 0.f;
        T* dst = out + (int64_t)b * p.out_row_stride + (int64_t)h * D + 4 * lane;
        dst[0] = from_f32<T>(o.x * inv);
        dst[1] = from_f32<T>(o.y * inv);
        dst[2] = from_f32<T>(o.z * inv);
        dst[3] = from_f32<T>(o.w * inv);
      }
      if (lane == 0) {
        const int old = atomicAdd(ticket, 1);
        if (old == S + G - 1) *ticket = 0;  // all S splits counted and all G rows combined: re-arm
      }
    }
    r = *rnext;
    consumer_sync();  // scratch and rnext are rewritten by the next row
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


__global__ void decode_workspace_init_kernel(uint32_t* sched) { sched[3] = kWorkspaceCookie; }

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

static int decode_ctas_per_sm() {
  static int v = 0;
  if (!v) {
    const char* e = tuning_env("MVB200_DECODE_CTAS_PER_SM");
    v = e ? atoi(e) : 2;
    if (v < 1) v = 1;
    if (v > 2) v = 2;
  }
  return v;
}

// MVB200_DECODE_REDUCE = fused | kernel: combine the split partials inside the persistent kernel (phase 2)
// or with decode_reduce_kernel launched behind it (PDL).
static bool decode_fused_reduce() {
  static const bool v = [] {
    const char* e = tuning_env("MVB200_DECODE_REDUCE");
    return !(e && e[0] == 'k');
  }();
  return v;
}

template <typename T, int D>
static int launch_fast(DecodeParams& p, const void* k_cache, const void* v_cache,
                       int block_size, cudaStream_t stream) {
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
  // ring | Q slots | barriers | unit info | per-warp max/sum | alignment slack
  const int smem = p.stages * (kStageBytes + kQSlotBytes) + p.stages * (16 + (int)sizeof(UnitInfo)) +
                   2 * kConsumerWarps * kMaxGroup * 4 + 1024;
  auto kern = paged_decode_kernel<T, D>;
  static SmemOptIn optin;
  rc = optin_dynamic_smem((const void*)kern, smem, &optin, "paged_decode");
  if (rc) return rc;
  // persistent grid: the CTAs that can be resident at once, or one per unit when there are fewer units.
  // "Can be resident" is asked of the runtime, not assumed: the fused split reduction waits on tickets owed by
  // other CTAs of the grid, so a CTA that is not co-resident with the waiters would never pay — a hang, not a
  // slowdown.  More stages (tuning build), register growth or another carve-out lower the answer below
  // decode_ctas_per_sm(); the grid follows it.
  static ResidentCtas resident;  // per instantiation: (device, smem bytes) -> CTAs per SM
  int per_sm = 0;
  rc = resident_ctas_per_sm((const void*)kern, kDecodeThreads, smem, &resident, "paged_decode", &per_sm);
  if (rc) return rc;
  if (per_sm > decode_ctas_per_sm()) per_sm = decode_ctas_per_sm();
  const int64_t capacity = (int64_t)per_sm * device_sm_count();
  int64_t ctas = capacity;
  if (ctas > p.total_units) ctas = p.total_units;
  // run-ahead pays where start-up latency shows: at most one unit per resident CTA (measured, run 7: with
  // several units per CTA — 512 x 1024-token sequences on one kv head — it costs 7 us per launch)
  if (p.total_units > capacity) p.early = 0;
  // a run-ahead producer would queue the unit's Q copy behind the tiles it fetched early: the consumers (which wait
  // for the preceding kernel at the top) read their Q fragments from global memory instead
  if (p.early && tuning_env("MVB200_DECODE_EARLY_QBULK") == nullptr) p.q_bulk = 0;
  const bool fused = p.num_splits > 1 && decode_fused_reduce();
  if (p.num_splits > 1 && !fused) p.tickets = nullptr;
  // one unit per CTA: nothing to balance, skip the counter round trips
  if (!fused && p.total_units <= ctas && tuning_env("MVB200_DECODE_DYNAMIC") == nullptr) p.sched = nullptr;
  cudaError_t le = launch_pdl(kern, dim3((unsigned)ctas), dim3(kDecodeThreads), (size_t)smem, stream, mk, mv, p);
  if (le != cudaSuccess)
    return fail(MVB200_ECUDA, "paged_decode_kernel: %s", cudaGetErrorString(le));
  rc = check_launch("paged_decode_kernel");
  if (rc || p.num_splits == 1 || fused) return rc;
  dim3 rgrid(p.num_heads, p.batch);
  le = launch_pdl(decode_reduce_kernel<T>, rgrid, dim3(128), 0, stream, (const float*)p.part_o,
                  (const float*)p.part_lse, (T*)p.out, p.out_row_stride, p.num_heads, D, p.num_splits);
  if (le != cudaSuccess) return fail(MVB200_ECUDA, "decode_reduce_kernel: %s", cudaGetErrorString(le));
  return check_launch("decode_reduce_kernel");
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
    const char* e = tuning_env("MVB200_DECODE_STAGES");
    v = e ? atoi(e) : 3;
    if (v < 2) v = 2;
    if (v > 6) v = 6;
  }
  return v;
}

}  // namespace mvb200

static size_t decode_partial_bytes(int32_t batch, int32_t num_heads, int32_t head_dim, int32_t num_splits) {
  if (num_splits <= 1 || batch <= 0) return 0;
  return (size_t)batch * num_heads * num_splits * ((size_t)head_dim + 1) * sizeof(float);
}

extern "C" size_t mvb200_decode_workspace_bytes(int32_t batch, int32_t num_heads,
                                                int32_t head_dim, int32_t num_splits) {
  if (batch <= 0) return 0;
  return decode_partial_bytes(batch, num_heads, head_dim, num_splits) + mvb200::kDecodeTailBytes;
}

extern "C" int mvb200_decode_workspace_init(void* workspace, size_t workspace_bytes, void* stream) {
  using namespace mvb200;
  if (!workspace || workspace_bytes < kDecodeTailBytes || ((uintptr_t)workspace & 15) || (workspace_bytes & 15))
    return fail(MVB200_EINVAL, "decode_workspace_init: need a 16-byte aligned buffer of at least %zu bytes "
                               "(a multiple of 16)", kDecodeTailBytes);
  cudaStream_t s = (cudaStream_t)stream;
  char* tail = (char*)workspace + workspace_bytes - kDecodeTailBytes;
  cudaError_t e = cudaMemsetAsync(tail, 0, kDecodeTailBytes, s);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "decode_workspace_init: %s", cudaGetErrorString(e));
  decode_workspace_init_kernel<<<1, 1, 0, s>>>((uint32_t*)(tail + (size_t)kDecodeMaxTickets * 4));
  return check_launch("decode_workspace_init_kernel");
}

extern "C" int32_t mvb200_decode_suggest_splits(int32_t batch, int32_t num_kv_heads,
                                                int32_t max_context, int32_t sm_count) {
  if (batch <= 0 || num_kv_heads <= 0) return 1;
  if (sm_count <= 0) sm_count = 148;
  const int64_t units = (int64_t)batch * num_kv_heads;
  const int tiles = max_context > 0 ? (max_context + 63) / 64 : 1;
  // The kernel is persistent (two resident CTAs per SM walk the (sequence, split, kv head) units, drawing
  // them from a device-side counter), so plenty of units need no split at all.  Fewer units than resident
  // CTAs: split until every CTA slot has a unit, keeping >= 4 tiles (128 KiB of K+V) per unit.
  // Measured on B200 (profiles/r02_decode_sweep_*): 256 units (one TP8 rank of cfg2) 44 us unsplit vs 48-51 us
  // with 2-4 splits; 4 units (1 x 32k) 21-22 us at 37-74 splits vs 28 us at 18.
  const int64_t slots = 2LL * sm_count;
  int64_t s = units >= sm_count ? 1 : slots / units;  // never more units than resident CTA slots: one wave
  static const int env_cap = [] {
    const char* e = mvb200::tuning_env("MVB200_DECODE_MAX_SPLITS");
    return e ? atoi(e) : 0;
  }();
  const int max_by_work = tiles / 4 > 0 ? tiles / 4 : 1;
  if (s > max_by_work) s = max_by_work;
  if (env_cap > 0 && s > env_cap) s = env_cap;
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
  return mvb200_paged_decode_ex(q, q_row_stride, k_cache, v_cache, block_tables, bt_row_stride, context_lens, out,
                                out_row_stride, workspace, workspace_bytes, batch, num_heads, num_kv_heads, head_dim,
                                block_size, num_blocks, max_blocks_per_seq, num_splits, scale, dtype, 0u, nullptr, stream);
}

extern "C" int mvb200_paged_decode_ex(const void* q, int64_t q_row_stride, const void* k_cache,
                                      const void* v_cache, const int32_t* block_tables,
                                      int64_t bt_row_stride, const int32_t* context_lens, void* out,
                                      int64_t out_row_stride, void* workspace, size_t workspace_bytes,
                                      int32_t batch, int32_t num_heads, int32_t num_kv_heads,
                                      int32_t head_dim, int32_t block_size, int64_t num_blocks,
                                      int32_t max_blocks_per_seq, int32_t num_splits, float scale,
                                      int32_t dtype, uint32_t flags, const mvb200_decode_store* store,
                                      void* stream) {
  using namespace mvb200;
  // Every argument check comes before the first write: a call that returns an error has not touched the cache
  // (the separate-launch path, store_kv_cache then attention, fails the same way: attention.py:237-243).
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
  const bool zero_width = max_blocks_per_seq <= 0 || num_blocks <= 0;
  if (!zero_width && (!k_cache || !v_cache || !block_tables))
    return fail(MVB200_EINVAL, "paged_decode: null cache or block table");
  if (store != nullptr && k_cache && v_cache && num_blocks > 0 &&
      (!store->key || !store->value || !store->slot_mapping))
    return fail(MVB200_EINVAL, "paged_decode: fused store with a null pointer");
  const int group = num_heads / num_kv_heads;
  const bool fast = !zero_width && eb == 2 && (head_dim == 64 || head_dim == 128) && block_size % 64 == 0 &&
                    group <= 16 && num_blocks < (1LL << 31) &&
                    ((uintptr_t)k_cache % 16 == 0) && ((uintptr_t)v_cache % 16 == 0) &&
                    ((uintptr_t)q % 4 == 0) && (q_row_stride % 2 == 0);
  if (!zero_width && !fast) {
    if (head_dim > 256) return fail(MVB200_EUNSUPPORTED, "paged_decode: head_dim %d > 256", head_dim);
    if (batch > 65535) return fail(MVB200_EUNSUPPORTED, "paged_decode: batch %d > 65535", batch);
  }
  if (store != nullptr && k_cache && v_cache && num_blocks > 0) {
    const bool fusable = fast && max_blocks_per_seq > 0 &&
                         ((uintptr_t)store->key % 16 == 0) && ((uintptr_t)store->value % 16 == 0) &&
                         store->key_row_stride % 8 == 0 && store->value_row_stride % 8 == 0 &&
                         tuning_env("MVB200_DECODE_NO_FUSED_STORE") == nullptr;
    if (!fusable) {
      const int rc = mvb200_store_kvcache(store->key, store->value, store->key_row_stride, store->value_row_stride,
                                          (void*)k_cache, (void*)v_cache, store->slot_mapping, batch,
                                          num_kv_heads * head_dim, num_blocks * (int64_t)block_size, dtype, stream);
      if (rc) return rc;
      store = nullptr;
    }
  } else {
    store = nullptr;
  }
  if (zero_width) {
    // zero-width block table (engine warm-up): zeros, like page_attention.py:62-65
    for (int b = 0; b < batch; ++b) {
      cudaError_t e = cudaMemsetAsync((char*)out + (int64_t)b * out_row_stride * eb, 0,
                                      (size_t)num_heads * head_dim * eb, s);
      if (e != cudaSuccess) return fail(MVB200_ECUDA, "paged_decode: memset: %s", cudaGetErrorString(e));
    }
    note_kernel("decode_zero");
    return MVB200_OK;
  }
  if (num_splits < 1) num_splits = 1;
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
    p.tickets = nullptr;
    p.sched = nullptr;
    p.batch = batch;
    const int64_t units = (int64_t)batch * num_kv_heads;
    // Without an (initialised) workspace, or beyond the ticket capacity, the kernel still works: no
    // splits, units dealt round-robin instead of drawn from the device-side counter.
    const bool have_tail = workspace && workspace_bytes >= kDecodeTailBytes && units <= kDecodeMaxTickets &&
                           ((uintptr_t)workspace & 15) == 0 && (workspace_bytes & 15) == 0;
    if (num_splits > 1) {
      const size_t need = mvb200_decode_workspace_bytes(batch, num_heads, head_dim, num_splits);
      if (!have_tail || workspace_bytes < need)
        return fail(MVB200_EINVAL, "paged_decode: workspace %zu B < %zu B needed for %d splits (or batch * "
                                   "kv heads > %d, or buffer not 16-byte aligned)",
                    workspace_bytes, need, num_splits, kDecodeMaxTickets);
      p.part_o = (float*)workspace;
      p.part_lse = p.part_o + (size_t)batch * num_heads * num_splits * head_dim;
    }
    if (have_tail) {
      char* tail = (char*)workspace + workspace_bytes - kDecodeTailBytes;
      p.tickets = (int32_t*)tail;
      p.sched = (int32_t*)(tail + (size_t)kDecodeMaxTickets * 4);
    }
    if (units * num_splits >= (1LL << 31)) return fail(MVB200_EINVAL, "paged_decode: too many work units");
    p.total_units = (int32_t)(units * num_splits);
    p.q_bulk = (((uintptr_t)q & 15) == 0 && (q_row_stride % 8) == 0) ? 1 : 0;
    static const bool no_early = tuning_env("MVB200_DECODE_NO_EARLY") != nullptr;
    // Run-ahead is requested here and confirmed in launch_fast, which knows how many CTAs are resident (it is
    // kept only with at most one unit per resident CTA).  The fused store does not stand in the way: the new rows
    // are read by a CONSUMER warp, after its own dependency wait (runs 31/32: one-launch step of the small
    // shapes -7 ... -11 %).
    p.early = ((flags & MVB200_DECODE_KV_PREFIX_STABLE) && pdl_enabled() && !no_early) ? 1 : 0;
    p.new_k = store ? store->key : nullptr;
    p.new_v = store ? store->value : nullptr;
    p.new_k_stride = store ? store->key_row_stride : 0;
    p.new_v_stride = store ? store->value_row_stride : 0;
    p.slot_mapping = store ? store->slot_mapping : nullptr;
    p.k_cache = (void*)k_cache;
    p.v_cache = (void*)v_cache;
    p.num_slots = num_blocks * (int64_t)block_size;
    if (store) note_kernel("store_fused_in_decode");
    note_kernel(num_splits > 1 ? "decode_tma_mma_fused_reduce" : "decode_tma_mma");
    if (dtype == MVB200_BF16) {
      return head_dim == 128 ? launch_fast<__nv_bfloat16, 128>(p, k_cache, v_cache, block_size, s)
                             : launch_fast<__nv_bfloat16, 64>(p, k_cache, v_cache, block_size, s);
    }
    return head_dim == 128 ? launch_fast<__half, 128>(p, k_cache, v_cache, block_size, s)
                           : launch_fast<__half, 64>(p, k_cache, v_cache, block_size, s);
  }
  if (store != nullptr) {  // the generic kernel does not store: separate launch
    const int rc = mvb200_store_kvcache(store->key, store->value, store->key_row_stride, store->value_row_stride,
                                        (void*)k_cache, (void*)v_cache, store->slot_mapping, batch,
                                        num_kv_heads * head_dim, num_blocks * (int64_t)block_size, dtype, stream);
    if (rc) return rc;
  }
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
