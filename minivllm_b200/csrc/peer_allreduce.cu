// peer_allreduce.cu — one-shot all-reduce over NVLink peer memory fused with the residual add and
// RMSNorm that follow it.
//
// Reference operator sequence being replaced, per decoder layer and per decode step:
//   RowParallelLinear.forward : dist.all_reduce(y)                      (models/layers/linear.py:314-319)
//   RMSNorm.forward(x, residual): x = x.float() + residual.float(); residual = x.to(dtype);
//                                 y = (x * rsqrt(mean(x^2) + eps)).to(dtype) * w     (layernorm.py:81-105)
// At decode the message is B x hidden elements (cfg3: 8 x 2560 bf16 = 40 KiB): NCCL's launch + protocol
// latency dominates and the two following elementwise kernels re-read what the collective just wrote.
// Here ONE kernel per rank
//   1. pushes its partial rows into a slot of every peer's receive buffer (posted 16-byte NVLink
//      stores, cudaIpc-mapped peer memory on an NVSwitch fabric: every peer is one hop away),
//   2. publishes a per-CTA epoch flag on every peer (st.release.sys) and waits for the peers' flags
//      (ld.acquire.sys) — no grid-wide or host synchronisation, CTA c of every rank owns the same rows,
//   3. sums the world_size slots from its own HBM/L2 in rank order in fp32 (every rank gets bit-identical
//      results), rounds once to the model dtype (the all-reduce output), adds the residual, writes the new
//      residual and the normalised row.
// (A two-shot variant for larger messages — reduce-scatter by rows, all-gather of the finished rows — is
// further down.)  The epoch lives in device memory and is bumped by the kernel itself, so the launch is CUDA-Graph
// capturable; receive buffers are double-buffered by epoch parity (a rank can only be one call ahead of
// a peer, see wait below).  Latency-bound by design: ~2 NVLink traversals per call.
#include "common.cuh"

namespace mvb200 {

constexpr int kMaxPeers = 8;
constexpr int kMaxCtas = 256;    // flag slots per source rank (one CTA per row up to 256 rows)
constexpr int kArThreads = 256;
constexpr int kMaxChunks = 4;    // 16-byte chunks per thread per row: hidden*sizeof(T) <= 256*16*4 = 16 KiB

struct PeerComm {
  char* data[kMaxPeers];          // receive buffers: [2 parities][world slots][slot_bytes]
  uint32_t* flags[kMaxPeers];     // [2 phases][kMaxPeers sources][kMaxCtas]
  uint32_t* epoch;                // local, [kMaxCtas]
  int64_t slot_bytes;
  int32_t rank, world;
};

MVB_DEV void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
MVB_DEV uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
MVB_DEV uint4 ld_cg(const void* p) {  // L2 only: the line was written by a peer through NVLink
  uint4 v;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

template <typename T>
struct Pack;  // 16 bytes of T <-> floats
template <>
struct Pack<float> {
  static constexpr int N = 4;
  MVB_DEV static void unpack(const uint4& r, float* f) {
    f[0] = __uint_as_float(r.x); f[1] = __uint_as_float(r.y); f[2] = __uint_as_float(r.z); f[3] = __uint_as_float(r.w);
  }
  MVB_DEV static uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
  }
};
template <>
struct Pack<__nv_bfloat16> {
  static constexpr int N = 8;
  MVB_DEV static void unpack(const uint4& r, float* f) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  }
  MVB_DEV static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 v = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&v);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};
template <>
struct Pack<__half> {
  static constexpr int N = 8;
  MVB_DEV static void unpack(const uint4& r, float* f) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 v = __half22float2(*reinterpret_cast<const __half2*>(&w[i]));
      f[2 * i] = v.x;
      f[2 * i + 1] = v.y;
    }
  }
  MVB_DEV static uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 v = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&v);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// ---- shared pieces ---------------------------------------------------------------------------------
// Spin until *flag has reached `epoch` (a peer may already be one call ahead).  A peer that never arrives
// is a protocol or launch-order bug: trap after ~20 s of SM clocks instead of hanging the GPU.
MVB_DEV void wait_flag(const uint32_t* flag, uint32_t epoch, int rank, int from, int cta, int phase) {
  const long long t0 = clock64();
  while ((int32_t)(ld_acquire_sys(flag) - epoch) < 0) {
    if (clock64() - t0 > 40000000000LL) {
      printf("mvb200: peer all-reduce timed out (rank %d waits for rank %d, cta %d, phase %d, epoch %u, saw %u)\n",
             rank, from, cta, phase, epoch, ld_acquire_sys(flag));
      __trap();
    }
  }
}

// One row, whole CTA: sum the `world` partials in rank order in fp32 (load(s, c) = 16-byte chunk c of rank s),
// round once to T (what dist.all_reduce returns), add the residual (kept in fp32 for the norm, written back
// rounded: layernorm.py:96-99), normalise.  emit(c, y16, res16) receives the packed outputs per chunk.
template <typename T, typename Load, typename Emit>
MVB_DEV void reduce_norm_row(Load load, Emit emit, int world, int nchunk, int hidden, float eps,
                             const uint4* res_row /* this row's residual or null */,
                             const uint4* res_pf /* prefetched residual chunks or null */, bool has_weight,
                             const uint4* w_pf, float* red) {
  constexpr int N = Pack<T>::N;
  const int tid = threadIdx.x;
  float vals[kMaxChunks][N];
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < kMaxChunks; ++it) {
    const int c = tid + it * kArThreads;
    if (c < nchunk) {
      float acc[N];
#pragma unroll
      for (int e = 0; e < N; ++e) acc[e] = 0.f;
      for (int s = 0; s < world; ++s) {
        float f[N];
        Pack<T>::unpack(load(s, c), f);
#pragma unroll
        for (int e = 0; e < N; ++e) acc[e] += f[e];
      }
      uint4 summed = Pack<T>::pack(acc);
      Pack<T>::unpack(summed, vals[it]);
      if (res_row) {
        float f[N];
        Pack<T>::unpack(res_pf ? res_pf[it] : res_row[c], f);
#pragma unroll
        for (int e = 0; e < N; ++e) vals[it][e] += f[e];
        summed = Pack<T>::pack(vals[it]);
      }
      if (!has_weight) emit(c, summed, summed);
#pragma unroll
      for (int e = 0; e < N; ++e) ss += vals[it][e] * vals[it][e];
    }
  }
  if (!has_weight) return;  // uniform across the CTA
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
  __syncthreads();  // red[] free again
  if ((tid & 31) == 0) red[tid >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
#pragma unroll
  for (int i = 0; i < kArThreads / 32; ++i) tot += red[i];
  const float inv = 1.f / sqrtf(tot / (float)hidden + eps);
#pragma unroll
  for (int it = 0; it < kMaxChunks; ++it) {
    const int c = tid + it * kArThreads;
    if (c < nchunk) {
      float w[N], o[N];
      Pack<T>::unpack(w_pf[it], w);
#pragma unroll
      for (int e = 0; e < N; ++e) o[e] = __fmul_rn(__fmul_rn(vals[it][e], inv), w[e]);
      emit(c, Pack<T>::pack(o), Pack<T>::pack(vals[it]));
    }
  }
}

MVB_DEV char* row_ptr(void* base, int64_t row, int64_t stride_bytes) {
  return reinterpret_cast<char*>(base) + row * stride_bytes;
}

// ---- one-shot: every rank pushes everything to everybody ------------------------------------------
// grid = G CTAs (the same on every rank); CTA c owns rows c, c+G, ... (row -> CTA is r % cap for every
// row count, so a row keeps its CTA, and with it its epoch parity, from call to call).
// x [rows, hidden] partial sums of this rank (row stride xs elements);
// residual (nullable) [rows, hidden] updated in place; weight (nullable) [hidden];
// y [rows, hidden]: weight ? RMSNorm(sum + residual) * weight : sum (+ residual).
template <typename T>
__global__ void __launch_bounds__(kArThreads)
peer_allreduce_kernel(PeerComm comm, const T* __restrict__ x, int64_t xs, T* __restrict__ residual,
                      int64_t rs, const T* __restrict__ weight, T* __restrict__ y, int64_t ys,
                      int32_t rows, int32_t hidden, float eps) {
  constexpr int N = Pack<T>::N;
  __shared__ float red[kArThreads / 32];
  __shared__ uint32_t s_epoch;
  const int cta = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
  const int nchunk = hidden / N;  // 16-byte chunks per row
  const int64_t row_bytes = (int64_t)hidden * sizeof(T);

  // x comes from the GEMM before, residual from the norm before that, and epoch[] from the previous
  // peer all-reduce (which may be the kernel right before this one): nothing is read before the wait.
  pdl_wait();
  pdl_launch_dependents();
  if (tid == 0) s_epoch = comm.epoch[cta] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const int64_t parity_off = (int64_t)(epoch & 1) * (comm.world + 2) * comm.slot_bytes;

  // 1. push my rows into slot [parity][rank] of every peer (self included: one code path, one sum order)
  for (int r = cta; r < rows; r += G) {
    const char* src = reinterpret_cast<const char*>(x + (int64_t)r * xs);
    for (int c = tid; c < nchunk; c += kArThreads) {
      const uint4 v = *reinterpret_cast<const uint4*>(src + (int64_t)c * 16);
      const int64_t off = parity_off + comm.rank * comm.slot_bytes + r * row_bytes + (int64_t)c * 16;
#pragma unroll 1
      for (int p = 0; p < comm.world; ++p) {
        const int peer = (comm.rank + p) % comm.world;  // spread the first stores over the links
        *reinterpret_cast<uint4*>(comm.data[peer] + off) = v;
      }
    }
  }
  // 2. publish + wait.  Completing call n+1 needs every peer's flag for n+1, which a peer publishes only
  //    after it finished call n entirely (stream order): so parity n&1 is never rewritten (in call n+2)
  //    while somebody still reads call n.  The release store is cumulative over the CTA barrier: every
  //    thread's pushes happen-before it.  While the flags travel, prefetch what step 3 needs locally.
  uint4 res_pf[kMaxChunks], w_pf[kMaxChunks];
#pragma unroll
  for (int it = 0; it < kMaxChunks; ++it) {
    const int c = tid + it * kArThreads;
    res_pf[it] = make_uint4(0, 0, 0, 0);
    w_pf[it] = make_uint4(0, 0, 0, 0);
    if (c < nchunk) {
      if (residual && cta < rows)
        res_pf[it] = *reinterpret_cast<const uint4*>(row_ptr(residual, cta, rs * sizeof(T)) + (int64_t)c * 16);
      if (weight) w_pf[it] = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(weight) + (int64_t)c * 16);
    }
  }
  __syncthreads();
  if (tid < comm.world) {
    st_release_sys(comm.flags[tid] + comm.rank * kMaxCtas + cta, epoch);
    wait_flag(comm.flags[comm.rank] + tid * kMaxCtas + cta, epoch, comm.rank, tid, cta, 1);
  }
  __syncthreads();
  if (tid == 0) comm.epoch[cta] = epoch;

  // 3. reduce + residual + norm, all local
  const char* base = comm.data[comm.rank] + parity_off;
  for (int r = cta; r < rows; r += G) {
    char* res_row = residual ? row_ptr(residual, r, rs * sizeof(T)) : nullptr;
    char* y_row = row_ptr(y, r, ys * sizeof(T));
    reduce_norm_row<T>(
        [&](int s, int c) { return ld_cg(base + s * comm.slot_bytes + r * row_bytes + (int64_t)c * 16); },
        [&](int c, const uint4& yv, const uint4& rv) {
          *reinterpret_cast<uint4*>(y_row + (int64_t)c * 16) = yv;
          if (res_row) *reinterpret_cast<uint4*>(res_row + (int64_t)c * 16) = rv;
        },
        comm.world, nchunk, hidden, eps, reinterpret_cast<const uint4*>(res_row), r == cta ? res_pf : nullptr,
        weight != nullptr, w_pf, red);
  }
}

// ---- two-shot: reduce-scatter by rows, all-gather of the normalised rows ---------------------------
// For messages past the latency-bound range.  CTA c of every rank handles rows c, c+G, ... (G a multiple
// of world) and those rows are OWNED by rank c % world:
//   non-owner: push my partial rows into the owner's slot [parity][rank]            (1/world of the
//              message per peer instead of all of it), release flag1 on the owner, wait for the owner's
//              flag2, copy the finished rows from my own Y (and R) region into y (and residual);
//   owner:     wait flag1 of every peer, reduce (own partial straight from x) + residual + norm, write
//              y/residual locally and push them into every peer's Y (/R) region, release flag2 everywhere.
// Per rank and call: (world-1)/world of the message in, the same out (twice that when the residual is
// gathered too) instead of (world-1) x the message with one-shot.  gather_residual == 0 leaves the residual
// stream row-sharded: each rank keeps only the rows it owns up to date (enough for a decoder whose residual
// is consumed by these kernels only; the first call of a forward pass starts from a replicated residual).
template <typename T>
__global__ void __launch_bounds__(kArThreads)
peer_allreduce_twoshot_kernel(PeerComm comm, const T* __restrict__ x, int64_t xs, T* __restrict__ residual,
                              int64_t rs, const T* __restrict__ weight, T* __restrict__ y, int64_t ys,
                              int32_t rows, int32_t hidden, float eps, int32_t gather_residual) {
  constexpr int N = Pack<T>::N;
  __shared__ float red[kArThreads / 32];
  __shared__ uint32_t s_epoch;
  const int cta = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
  const int nchunk = hidden / N;
  const int64_t row_bytes = (int64_t)hidden * sizeof(T);
  const int owner = cta % comm.world;
  const bool mine = owner == comm.rank;
  const bool push_res = residual != nullptr && gather_residual != 0;

  pdl_wait();
  pdl_launch_dependents();
  if (tid == 0) s_epoch = comm.epoch[cta] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const int64_t parity_off = (int64_t)(epoch & 1) * (comm.world + 2) * comm.slot_bytes;
  const int64_t y_off = parity_off + (int64_t)comm.world * comm.slot_bytes;  // Y region, R region follows

  if (!mine) {
    for (int r = cta; r < rows; r += G) {
      const char* src = reinterpret_cast<const char*>(x + (int64_t)r * xs);
      char* dst = comm.data[owner] + parity_off + comm.rank * comm.slot_bytes + r * row_bytes;
      for (int c = tid; c < nchunk; c += kArThreads)
        *reinterpret_cast<uint4*>(dst + (int64_t)c * 16) = *reinterpret_cast<const uint4*>(src + (int64_t)c * 16);
    }
    __syncthreads();
    if (tid == 0) {
      st_release_sys(comm.flags[owner] + comm.rank * kMaxCtas + cta, epoch);
      wait_flag(comm.flags[comm.rank] + (kMaxPeers + owner) * kMaxCtas + cta, epoch, comm.rank, owner, cta, 2);
    }
    __syncthreads();
    if (tid == 0) comm.epoch[cta] = epoch;
    const char* ybase = comm.data[comm.rank] + y_off;
    for (int r = cta; r < rows; r += G) {
      for (int c = tid; c < nchunk; c += kArThreads) {
        *reinterpret_cast<uint4*>(row_ptr(y, r, ys * sizeof(T)) + (int64_t)c * 16) =
            ld_cg(ybase + r * row_bytes + (int64_t)c * 16);
        if (push_res)
          *reinterpret_cast<uint4*>(row_ptr(residual, r, rs * sizeof(T)) + (int64_t)c * 16) =
              ld_cg(ybase + comm.slot_bytes + r * row_bytes + (int64_t)c * 16);
      }
    }
    return;
  }

  // owner
  uint4 w_pf[kMaxChunks];
#pragma unroll
  for (int it = 0; it < kMaxChunks; ++it) {
    const int c = tid + it * kArThreads;
    w_pf[it] = make_uint4(0, 0, 0, 0);
    if (c < nchunk && weight)
      w_pf[it] = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(weight) + (int64_t)c * 16);
  }
  if (tid < comm.world && tid != comm.rank)
    wait_flag(comm.flags[comm.rank] + tid * kMaxCtas + cta, epoch, comm.rank, tid, cta, 1);
  __syncthreads();
  const char* base = comm.data[comm.rank] + parity_off;
  for (int r = cta; r < rows; r += G) {
    const char* x_row = reinterpret_cast<const char*>(x + (int64_t)r * xs);
    char* res_row = residual ? row_ptr(residual, r, rs * sizeof(T)) : nullptr;
    char* y_row = row_ptr(y, r, ys * sizeof(T));
    reduce_norm_row<T>(
        [&](int s, int c) {
          return s == comm.rank ? *reinterpret_cast<const uint4*>(x_row + (int64_t)c * 16)
                                : ld_cg(base + s * comm.slot_bytes + r * row_bytes + (int64_t)c * 16);
        },
        [&](int c, const uint4& yv, const uint4& rv) {
          *reinterpret_cast<uint4*>(y_row + (int64_t)c * 16) = yv;
          if (res_row) *reinterpret_cast<uint4*>(res_row + (int64_t)c * 16) = rv;
          const int64_t off = y_off + r * row_bytes + (int64_t)c * 16;
#pragma unroll 1
          for (int p = 1; p < comm.world; ++p) {
            const int peer = (comm.rank + p) % comm.world;
            *reinterpret_cast<uint4*>(comm.data[peer] + off) = yv;
            if (push_res) *reinterpret_cast<uint4*>(comm.data[peer] + off + comm.slot_bytes) = rv;
          }
        },
        comm.world, nchunk, hidden, eps, reinterpret_cast<const uint4*>(res_row), nullptr, weight != nullptr,
        w_pf, red);
  }
  __syncthreads();
  if (tid < comm.world && tid != comm.rank)
    st_release_sys(comm.flags[tid] + (kMaxPeers + comm.rank) * kMaxCtas + cta, epoch);
  if (tid == 0) comm.epoch[cta] = epoch;
}

// ---- NVLS: the two-shot schedule with the switch doing the sum and the broadcast ------------------------
// Same ownership as the two-shot kernel (CTA c of every rank handles rows c, c+G, ...; rank c % world owns
// them), but the data lives in a SYMMETRIC buffer that is also mapped as an NVLink multicast object
// (torch.distributed._symmetric_memory is the plumbing: minivllm_b200/comm.py):
//   every rank : copy my partial rows into the X region of my own buffer (local stores), release flag1 on
//                the owner;
//   owner      : wait flag1 of every peer, multimem.ld_reduce the rows from the multicast address — the
//                NVSwitch sums the world copies in fp32 and returns one (one 1/world-sized read per rank
//                instead of world slots) — add the residual, normalise, multimem.st the finished rows into
//                everybody's Y (/R) region (one store, the switch replicates it), release flag2 everywhere;
//   non-owner  : wait flag2, copy the rows from its own Y (/R) region into y (/residual).
// Egress per rank ~ 1x the message instead of 1.75x (world 8) and 1/world of the store instructions.
// The sum order is the switch's: results equal the rank-order fp32 sum whenever that sum is exact in fp32
// (always for the few-term bf16 sums here, up to the final rounding).
struct NvlsComm {
  char* mc;             // multicast mapping of the symmetric buffer: [2 parities][X | Y | R][slot_bytes]
  char* local;          // this rank's own mapping of the same buffer
  int64_t slot_bytes;
};

template <typename T>
MVB_DEV uint4 multimem_ld_reduce(const char* mc_addr);
template <>
MVB_DEV uint4 multimem_ld_reduce<__nv_bfloat16>(const char* a) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a) : "memory");
  return v;
}
template <>
MVB_DEV uint4 multimem_ld_reduce<__half>(const char* a) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a) : "memory");
  return v;
}
template <>
MVB_DEV uint4 multimem_ld_reduce<float>(const char* a) {
  float4 f;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(f.x), "=f"(f.y), "=f"(f.z), "=f"(f.w) : "l"(a) : "memory");
  return make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z), __float_as_uint(f.w));
}
MVB_DEV void multimem_st(char* mc_addr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc_addr), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}

template <typename T>
__global__ void __launch_bounds__(kArThreads)
nvls_allreduce_kernel(PeerComm comm, NvlsComm nv, const T* __restrict__ x, int64_t xs, T* __restrict__ residual,
                      int64_t rs, const T* __restrict__ weight, T* __restrict__ y, int64_t ys, int32_t rows,
                      int32_t hidden, float eps, int32_t gather_residual) {
  constexpr int N = Pack<T>::N;
  __shared__ float red[kArThreads / 32];
  __shared__ uint32_t s_epoch;
  const int cta = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
  const int nchunk = hidden / N;
  const int64_t row_bytes = (int64_t)hidden * sizeof(T);
  const int owner = cta % comm.world;
  const bool mine = owner == comm.rank;
  const bool push_res = residual != nullptr && gather_residual != 0;

  pdl_wait();
  pdl_launch_dependents();
  if (tid == 0) s_epoch = comm.epoch[cta] + 1;
  __syncthreads();
  const uint32_t epoch = s_epoch;
  const int64_t x_off = (int64_t)(epoch & 1) * 3 * nv.slot_bytes;  // X region; Y and R follow
  const int64_t y_off = x_off + nv.slot_bytes;

  // 1. my partial rows -> my own X region (the owner reads them through the switch)
  for (int r = cta; r < rows; r += G) {
    const char* src = reinterpret_cast<const char*>(x + (int64_t)r * xs);
    char* dst = nv.local + x_off + r * row_bytes;
    for (int c = tid; c < nchunk; c += kArThreads)
      *reinterpret_cast<uint4*>(dst + (int64_t)c * 16) = *reinterpret_cast<const uint4*>(src + (int64_t)c * 16);
  }
  __syncthreads();
  if (!mine) {
    if (tid == 0) {
      st_release_sys(comm.flags[owner] + comm.rank * kMaxCtas + cta, epoch);
      wait_flag(comm.flags[comm.rank] + (kMaxPeers + owner) * kMaxCtas + cta, epoch, comm.rank, owner, cta, 2);
    }
    __syncthreads();
    if (tid == 0) comm.epoch[cta] = epoch;
    const char* ybase = nv.local + y_off;
    for (int r = cta; r < rows; r += G) {
      for (int c = tid; c < nchunk; c += kArThreads) {
        *reinterpret_cast<uint4*>(row_ptr(y, r, ys * sizeof(T)) + (int64_t)c * 16) =
            ld_cg(ybase + r * row_bytes + (int64_t)c * 16);
        if (push_res)
          *reinterpret_cast<uint4*>(row_ptr(residual, r, rs * sizeof(T)) + (int64_t)c * 16) =
              ld_cg(ybase + nv.slot_bytes + r * row_bytes + (int64_t)c * 16);
      }
    }
    return;
  }

  // owner
  uint4 w_pf[kMaxChunks];
#pragma unroll
  for (int it = 0; it < kMaxChunks; ++it) {
    const int c = tid + it * kArThreads;
    w_pf[it] = make_uint4(0, 0, 0, 0);
    if (c < nchunk && weight)
      w_pf[it] = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(weight) + (int64_t)c * 16);
  }
  __threadfence_system();  // my own X rows are read back through the switch as well
  if (tid < comm.world && tid != comm.rank)
    wait_flag(comm.flags[comm.rank] + tid * kMaxCtas + cta, epoch, comm.rank, tid, cta, 1);
  __syncthreads();
  for (int r = cta; r < rows; r += G) {
    char* res_row = residual ? row_ptr(residual, r, rs * sizeof(T)) : nullptr;
    char* y_row = row_ptr(y, r, ys * sizeof(T));
    reduce_norm_row<T>(
        [&](int, int c) { return multimem_ld_reduce<T>(nv.mc + x_off + r * row_bytes + (int64_t)c * 16); },
        [&](int c, const uint4& yv, const uint4& rv) {
          *reinterpret_cast<uint4*>(y_row + (int64_t)c * 16) = yv;
          if (res_row) *reinterpret_cast<uint4*>(res_row + (int64_t)c * 16) = rv;
          char* dst = nv.mc + y_off + r * row_bytes + (int64_t)c * 16;
          multimem_st(dst, yv);
          if (push_res) multimem_st(dst + nv.slot_bytes, rv);
        },
        1, nchunk, hidden, eps, reinterpret_cast<const uint4*>(res_row), nullptr, weight != nullptr, w_pf, red);
  }
  __threadfence_system();  // the multicast stores are performed at every peer before the flags below
  __syncthreads();
  if (tid < comm.world && tid != comm.rank)
    st_release_sys(comm.flags[tid] + (kMaxPeers + comm.rank) * kMaxCtas + cta, epoch);
  if (tid == 0) comm.epoch[cta] = epoch;
}

template <typename T>
static int launch_ar(const PeerComm& comm, const void* x, int64_t xs, void* residual, int64_t rs,
                     const void* weight, void* y, int64_t ys, int32_t rows, int32_t hidden, float eps,
                     int32_t ctas, int32_t algo, int32_t gather_residual, cudaStream_t s) {
  cudaError_t e;
  if (algo == 2)
    e = launch_pdl(peer_allreduce_twoshot_kernel<T>, dim3(ctas), dim3(kArThreads), 0, s, comm, (const T*)x, xs,
                   (T*)residual, rs, (const T*)weight, (T*)y, ys, rows, hidden, eps, gather_residual);
  else
    e = launch_pdl(peer_allreduce_kernel<T>, dim3(ctas), dim3(kArThreads), 0, s, comm, (const T*)x, xs,
                   (T*)residual, rs, (const T*)weight, (T*)y, ys, rows, hidden, eps);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "peer_allreduce: %s", cudaGetErrorString(e));
  return check_launch("peer_allreduce");
}

}  // namespace mvb200

extern "C" {

int mvb200_peer_flag_bytes(void) { return 2 * mvb200::kMaxPeers * mvb200::kMaxCtas * (int)sizeof(uint32_t); }
int mvb200_peer_epoch_bytes(void) { return mvb200::kMaxCtas * (int)sizeof(uint32_t); }
int mvb200_peer_max_ctas(void) { return mvb200::kMaxCtas; }

int mvb200_peer_alloc(size_t bytes, void** ptr) {
  using namespace mvb200;
  if (!ptr || bytes == 0) return fail(MVB200_EINVAL, "peer_alloc: bad arguments");
  cudaError_t e = cudaMalloc(ptr, bytes);
  if (e == cudaSuccess) e = cudaMemset(*ptr, 0, bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "peer_alloc(%zu): %s", bytes, cudaGetErrorString(e));
  return MVB200_OK;
}

int mvb200_peer_free(void* ptr) {
  using namespace mvb200;
  cudaError_t e = cudaFree(ptr);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "peer_free: %s", cudaGetErrorString(e));
  return MVB200_OK;
}

int mvb200_peer_export(void* ptr, void* handle64) {
  using namespace mvb200;
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
  if (!ptr || !handle64) return fail(MVB200_EINVAL, "peer_export: null pointer");
  cudaError_t e = cudaIpcGetMemHandle(reinterpret_cast<cudaIpcMemHandle_t*>(handle64), ptr);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "cudaIpcGetMemHandle: %s", cudaGetErrorString(e));
  return MVB200_OK;
}

int mvb200_peer_open(const void* handle64, void** ptr) {
  using namespace mvb200;
  if (!ptr || !handle64) return fail(MVB200_EINVAL, "peer_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  cudaError_t e = cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e));
  return MVB200_OK;
}

int mvb200_peer_close(void* ptr) {
  using namespace mvb200;
  cudaError_t e = cudaIpcCloseMemHandle(ptr);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "cudaIpcCloseMemHandle: %s", cudaGetErrorString(e));
  return MVB200_OK;
}

int mvb200_peer_allreduce_add_rms_norm(void* const* data_ptrs, void* const* flag_ptrs, void* epoch,
                                       int64_t slot_bytes, int32_t rank, int32_t world, const void* x,
                                       int64_t x_row_stride, void* residual, int64_t residual_row_stride,
                                       const void* weight, void* y, int64_t y_row_stride, int32_t rows,
                                       int32_t hidden, float eps, int32_t dtype, int32_t max_ctas,
                                       int32_t algo, int32_t gather_residual, void* stream) {
  using namespace mvb200;
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
    return fail(MVB200_EINVAL, "peer_allreduce: bad rank/world (%d/%d, at most %d peers)", rank, world, kMaxPeers);
  if (rows < 0 || hidden <= 0) return fail(MVB200_EINVAL, "peer_allreduce: bad sizes (%d x %d)", rows, hidden);
  if (rows == 0) return MVB200_OK;
  const int eb = dtype == MVB200_F32 ? 4 : ((dtype == MVB200_F16 || dtype == MVB200_BF16) ? 2 : 0);
  if (!eb) return fail(MVB200_EINVAL, "peer_allreduce: unknown dtype %d", dtype);
  const int64_t row_bytes = (int64_t)hidden * eb;
  if (row_bytes % 16 || row_bytes > (int64_t)kArThreads * 16 * kMaxChunks)
    return fail(MVB200_EUNSUPPORTED, "peer_allreduce: row of %lld bytes (need a multiple of 16, at most %d)",
                (long long)row_bytes, kArThreads * 16 * kMaxChunks);
  if ((int64_t)rows * row_bytes > slot_bytes)
    return fail(MVB200_EINVAL, "peer_allreduce: %d rows of %lld bytes exceed the %lld-byte slot", rows,
                (long long)row_bytes, (long long)slot_bytes);
  if (!data_ptrs || !flag_ptrs || !epoch || !x || !y) return fail(MVB200_EINVAL, "peer_allreduce: null pointer");
  const uintptr_t mix = (uintptr_t)x | (uintptr_t)y | (uintptr_t)residual | (uintptr_t)weight |
                        (uintptr_t)(x_row_stride * eb) | (uintptr_t)(y_row_stride * eb) |
                        (uintptr_t)(residual ? residual_row_stride * eb : 0) | (uintptr_t)slot_bytes;
  if (mix & 15) return fail(MVB200_EINVAL, "peer_allreduce: pointers, strides and slot size must be 16-byte aligned");
  PeerComm comm;
  for (int i = 0; i < kMaxPeers; ++i) {
    comm.data[i] = i < world ? (char*)data_ptrs[i] : nullptr;
    comm.flags[i] = i < world ? (uint32_t*)flag_ptrs[i] : nullptr;
    if (i < world && (!comm.data[i] || !comm.flags[i])) return fail(MVB200_EINVAL, "peer_allreduce: null peer pointer %d", i);
  }
  comm.epoch = (uint32_t*)epoch;
  comm.slot_bytes = slot_bytes;
  comm.rank = rank;
  comm.world = world;
  if (algo != 1 && algo != 2) return fail(MVB200_EINVAL, "peer_allreduce: algo must be 1 (one-shot) or 2 (two-shot)");
  int32_t cap = (max_ctas > 0 && max_ctas < kMaxCtas) ? max_ctas : kMaxCtas;  // same on every rank
  cap -= cap % world;  // row r is always handled by CTA r % cap, whose owner (two-shot) is rank (r % cap) % world
  if (cap < world) return fail(MVB200_EINVAL, "peer_allreduce: max_ctas %d is smaller than the world size %d", max_ctas, world);
  const int32_t want = algo == 2 ? (rows + world - 1) / world * world : rows;
  const int32_t ctas = want < cap ? want : cap;
  note_kernel("peer_allreduce");
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case MVB200_BF16:
      return launch_ar<__nv_bfloat16>(comm, x, x_row_stride, residual, residual_row_stride, weight, y, y_row_stride, rows, hidden, eps, ctas, algo, gather_residual, s);
    case MVB200_F16:
      return launch_ar<__half>(comm, x, x_row_stride, residual, residual_row_stride, weight, y, y_row_stride, rows, hidden, eps, ctas, algo, gather_residual, s);
    default:
      return launch_ar<float>(comm, x, x_row_stride, residual, residual_row_stride, weight, y, y_row_stride, rows, hidden, eps, ctas, algo, gather_residual, s);
  }
}

int mvb200_nvls_allreduce_add_rms_norm(void* multicast_base, void* local_base, int64_t slot_bytes,
                                       void* const* flag_ptrs, void* epoch, int32_t rank, int32_t world,
                                       const void* x, int64_t x_row_stride, void* residual,
                                       int64_t residual_row_stride, const void* weight, void* y,
                                       int64_t y_row_stride, int32_t rows, int32_t hidden, float eps, int32_t dtype,
                                       int32_t max_ctas, int32_t gather_residual, void* stream) {
  using namespace mvb200;
  if (world < 2 || world > kMaxPeers || rank < 0 || rank >= world)
    return fail(MVB200_EINVAL, "nvls_allreduce: bad rank/world (%d/%d, 2..%d peers)", rank, world, kMaxPeers);
  if (rows < 0 || hidden <= 0) return fail(MVB200_EINVAL, "nvls_allreduce: bad sizes (%d x %d)", rows, hidden);
  if (rows == 0) return MVB200_OK;
  const int eb = dtype == MVB200_F32 ? 4 : ((dtype == MVB200_F16 || dtype == MVB200_BF16) ? 2 : 0);
  if (!eb) return fail(MVB200_EINVAL, "nvls_allreduce: unknown dtype %d", dtype);
  const int64_t row_bytes = (int64_t)hidden * eb;
  if (row_bytes % 16 || row_bytes > (int64_t)kArThreads * 16 * kMaxChunks)
    return fail(MVB200_EUNSUPPORTED, "nvls_allreduce: row of %lld bytes (need a multiple of 16, at most %d)",
                (long long)row_bytes, kArThreads * 16 * kMaxChunks);
  if ((int64_t)rows * row_bytes > slot_bytes)
    return fail(MVB200_EINVAL, "nvls_allreduce: %d rows of %lld bytes exceed the %lld-byte slot", rows,
                (long long)row_bytes, (long long)slot_bytes);
  if (!multicast_base || !local_base || !flag_ptrs || !epoch || !x || !y)
    return fail(MVB200_EINVAL, "nvls_allreduce: null pointer");
  const uintptr_t mix = (uintptr_t)x | (uintptr_t)y | (uintptr_t)residual | (uintptr_t)weight |
                        (uintptr_t)multicast_base | (uintptr_t)local_base | (uintptr_t)(x_row_stride * eb) |
                        (uintptr_t)(y_row_stride * eb) | (uintptr_t)(residual ? residual_row_stride * eb : 0) |
                        (uintptr_t)slot_bytes;
  if (mix & 15) return fail(MVB200_EINVAL, "nvls_allreduce: pointers, strides and slot size must be 16-byte aligned");
  PeerComm comm;
  for (int i = 0; i < kMaxPeers; ++i) {
    comm.data[i] = nullptr;
    comm.flags[i] = i < world ? (uint32_t*)flag_ptrs[i] : nullptr;
    if (i < world && !comm.flags[i]) return fail(MVB200_EINVAL, "nvls_allreduce: null peer flag pointer %d", i);
  }
  comm.epoch = (uint32_t*)epoch;
  comm.slot_bytes = slot_bytes;
  comm.rank = rank;
  comm.world = world;
  NvlsComm nv{(char*)multicast_base, (char*)local_base, slot_bytes};
  int32_t cap = (max_ctas > 0 && max_ctas < kMaxCtas) ? max_ctas : kMaxCtas;
  cap -= cap % world;
  if (cap < world) return fail(MVB200_EINVAL, "nvls_allreduce: max_ctas %d is smaller than the world size %d", max_ctas, world);
  const int32_t want = (rows + world - 1) / world * world;
  const int32_t ctas = want < cap ? want : cap;
  note_kernel("nvls_allreduce");
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t e;
#define MVB_NVLS_LAUNCH(T)                                                                                          \
  e = launch_pdl(nvls_allreduce_kernel<T>, dim3(ctas), dim3(kArThreads), 0, s, comm, nv, (const T*)x, x_row_stride,  \
                 (T*)residual, residual_row_stride, (const T*)weight, (T*)y, y_row_stride, rows, hidden, eps,        \
                 gather_residual)
  if (dtype == MVB200_BF16) MVB_NVLS_LAUNCH(__nv_bfloat16);
  else if (dtype == MVB200_F16) MVB_NVLS_LAUNCH(__half);
  else MVB_NVLS_LAUNCH(float);
#undef MVB_NVLS_LAUNCH
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "nvls_allreduce: %s", cudaGetErrorString(e));
  return check_launch("nvls_allreduce");
}

}  // extern "C"
