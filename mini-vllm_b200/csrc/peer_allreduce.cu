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
// The epoch lives in device memory and is bumped by the kernel itself, so the launch is CUDA-Graph
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
  uint32_t* flags[kMaxPeers];     // [world sources][kMaxCtas]
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

// grid = G CTAs (G <= kMaxCtas, the same on every rank); CTA c owns rows c, c+G, ...
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
  const int64_t parity_off = (int64_t)(epoch & 1) * comm.world * comm.slot_bytes;

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
  // 2. publish + wait.  A peer can start call n+1 (overwriting parity (n+1)&1) only after it saw my flag
  //    for call n+1... which I set after finishing call n, so parity n&1 is not rewritten before call n+2,
  //    and a peer enters call n+2 only after I published call n+1, i.e. after I finished reading call n.
  //    The release store is cumulative over the CTA barrier: every thread's pushes happen-before it.
  //    While the flags travel, prefetch what phase 3 needs from local memory (first row's residual, weight).
  uint4 res_pf[kMaxChunks], w_pf[kMaxChunks];
#pragma unroll
  for (int it = 0; it < kMaxChunks; ++it) {
    const int c = tid + it * kArThreads;
    res_pf[it] = make_uint4(0, 0, 0, 0);
    w_pf[it] = make_uint4(0, 0, 0, 0);
    if (c < nchunk) {
      if (residual && cta < rows)
        res_pf[it] = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(residual + (int64_t)cta * rs) + (int64_t)c * 16);
      if (weight) w_pf[it] = *reinterpret_cast<const uint4*>(reinterpret_cast<const char*>(weight) + (int64_t)c * 16);
    }
  }
  __syncthreads();
  if (tid < comm.world) {
    st_release_sys(comm.flags[tid] + comm.rank * kMaxCtas + cta, epoch);
    const uint32_t* mine = comm.flags[comm.rank] + tid * kMaxCtas + cta;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (clock64() - t0 > 40000000000LL) {  // ~20 s of SM clocks: a peer never arrived
        printf("mvb200: peer all-reduce timed out (rank %d waits for rank %d, cta %d, epoch %u, saw %u)\n",
               comm.rank, tid, cta, epoch, ld_acquire_sys(mine));
        __trap();
      }
    }
  }
  __syncthreads();
  if (tid == 0) comm.epoch[cta] = epoch;

  // 3. reduce in rank order, round to T (the all-reduce result), add residual, normalise
  const char* base = comm.data[comm.rank] + parity_off;
  for (int r = cta; r < rows; r += G) {
    float vals[kMaxChunks][N];
    float ss = 0.f;
#pragma unroll
    for (int it = 0; it < kMaxChunks; ++it) {
      const int c = tid + it * kArThreads;
      if (c < nchunk) {
        float acc[N];
#pragma unroll
        for (int e = 0; e < N; ++e) acc[e] = 0.f;
        for (int s = 0; s < comm.world; ++s) {
          float f[N];
          Pack<T>::unpack(ld_cg(base + s * comm.slot_bytes + r * row_bytes + (int64_t)c * 16), f);
#pragma unroll
          for (int e = 0; e < N; ++e) acc[e] += f[e];
        }
        uint4 summed = Pack<T>::pack(acc);  // dist.all_reduce returns the model dtype
        Pack<T>::unpack(summed, vals[it]);
        if (residual) {
          uint4* rp = reinterpret_cast<uint4*>(reinterpret_cast<char*>(residual + (int64_t)r * rs) + (int64_t)c * 16);
          float f[N];
          Pack<T>::unpack(r == cta ? res_pf[it] : *rp, f);
#pragma unroll
          for (int e = 0; e < N; ++e) vals[it][e] += f[e];
          summed = Pack<T>::pack(vals[it]);
          *rp = summed;  // residual = (x + residual).to(dtype); the norm uses the fp32 value (layernorm.py:96-99)
        }
        if (!weight)
          *reinterpret_cast<uint4*>(reinterpret_cast<char*>(y + (int64_t)r * ys) + (int64_t)c * 16) = summed;
#pragma unroll
        for (int e = 0; e < N; ++e) ss += vals[it][e] * vals[it][e];
      }
    }
    if (!weight) continue;  // uniform across the CTA
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
        *reinterpret_cast<uint4*>(reinterpret_cast<char*>(y + (int64_t)r * ys) + (int64_t)c * 16) = Pack<T>::pack(o);
      }
    }
  }
}

template <typename T>
static int launch_ar(const PeerComm& comm, const void* x, int64_t xs, void* residual, int64_t rs,
                     const void* weight, void* y, int64_t ys, int32_t rows, int32_t hidden, float eps,
                     int32_t ctas, cudaStream_t s) {
  cudaError_t e = launch_pdl(peer_allreduce_kernel<T>, dim3(ctas), dim3(kArThreads), 0, s, comm,
                             (const T*)x, xs, (T*)residual, rs, (const T*)weight, (T*)y, ys, rows, hidden, eps);
  if (e != cudaSuccess) return fail(MVB200_ECUDA, "peer_allreduce: %s", cudaGetErrorString(e));
  return check_launch("peer_allreduce");
}

}  // namespace mvb200

extern "C" {

int mvb200_peer_flag_bytes(void) { return mvb200::kMaxPeers * mvb200::kMaxCtas * (int)sizeof(uint32_t); }
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
                                       void* stream) {
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
  const int32_t cap = (max_ctas > 0 && max_ctas < kMaxCtas) ? max_ctas : kMaxCtas;  // same on every rank
  const int32_t ctas = rows < cap ? rows : cap;
  note_kernel("peer_allreduce");
  cudaStream_t s = (cudaStream_t)stream;
  switch (dtype) {
    case MVB200_BF16:
      return launch_ar<__nv_bfloat16>(comm, x, x_row_stride, residual, residual_row_stride, weight, y, y_row_stride, rows, hidden, eps, ctas, s);
    case MVB200_F16:
      return launch_ar<__half>(comm, x, x_row_stride, residual, residual_row_stride, weight, y, y_row_stride, rows, hidden, eps, ctas, s);
    default:
      return launch_ar<float>(comm, x, x_row_stride, residual, residual_row_stride, weight, y, y_row_stride, rows, hidden, eps, ctas, s);
  }
}

}  // extern "C"
