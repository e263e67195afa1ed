// common.cuh — sm_100a device helpers shared by the mvb200 kernels (PTX wrappers, error plumbing).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/mvb200.h"

namespace mvb200 {

// ---------------------------------------------------------------------------------------------
// host-side error plumbing (per-thread message + name of the last dispatched kernel family)
// ---------------------------------------------------------------------------------------------
int fail(int code, const char* fmt, ...);
void note_kernel(const char* name);
int check_launch(const char* what);

// Encode a tiled TMA descriptor (driver entry point resolved at run time; no libcuda link).
// dims/strides innermost first; strides are bytes for dims 1..rank-1.
int make_tensor_map(CUtensorMap* map, const void* base, int elem_bytes, int rank,
                    const uint64_t* dims, const uint64_t* strides_bytes, const uint32_t* box,
                    bool swizzle128);

// Launch with programmatic stream serialization (PDL): the kernel may start while its predecessor in
// the stream drains; it must execute pdl_wait() before it touches anything a predecessor wrote (or
// still reads).  Capturable into CUDA graphs.  MVB200_NO_PDL=1 falls back to a plain launch.
bool pdl_enabled();
// largest dynamic-smem size a kernel has been opted into, per device (one static instance per kernel)
struct SmemOptIn {
  int bytes[64];
};
int optin_dynamic_smem(const void* kernel, int bytes, SmemOptIn* cache, const char* what);
int device_sm_count();  // SMs of the current device (cached per device)
// CTAs of `kernel` (threads, dynamic smem) the runtime says fit one SM at a time, per device (one static instance
// per kernel).  Persistent kernels whose CTAs wait for each other size their grid from this, never from a guess.
struct ResidentCtas {
  int smem[64];
  int ctas[64];
};
int resident_ctas_per_sm(const void* kernel, int threads, int smem_bytes, ResidentCtas* cache, const char* what,
                         int* out);
// A/B knobs (MVB200_DECODE_*, MVB200_ROPE_PAIR, ...) exist in the tuning build only (python minivllm_b200/build.py
// --trace -> libmvb200_trace.so, selected with MVB200_LIB): the product library carries the measured defaults.
inline const char* tuning_env(const char* name) {
#ifdef MVB200_TUNING
  return getenv(name);
#else
  (void)name;
  return nullptr;
#endif
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---------------------------------------------------------------------------------------------
// device helpers
// ---------------------------------------------------------------------------------------------
#define MVB_DEV __device__ __forceinline__

// ---- programmatic dependent launch ----
MVB_DEV void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
MVB_DEV void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

MVB_DEV uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---- mbarrier ----
MVB_DEV void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
MVB_DEV void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
MVB_DEV void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
MVB_DEV void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
MVB_DEV bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Non-blocking probe (no suspend): for a role that polls several barriers.
MVB_DEV bool mbar_test(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Spin on try_wait (which itself sleeps in hardware).  A barrier that never flips is a protocol bug:
// trap after ~2 s of SM clocks instead of hanging the device.
MVB_DEV void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try_wait(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) {
      printf("mvb200: mbarrier wait timed out (block %d,%d,%d thread %d bar 0x%x parity %u)\n",
             blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x, bar, parity);
      __trap();
    }
  }
}

// ---- proxy fences ----
MVB_DEV void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---- TMA (cp.async.bulk.tensor), completion on an mbarrier ----
MVB_DEV void tma_prefetch_desc(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(map)) : "memory");
}
MVB_DEV void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                         int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
MVB_DEV void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1,
                         int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// ---- warp-level tensor ops used by the decode kernel ----
MVB_DEV void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
MVB_DEV void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}

template <typename T>
struct Mma16816;
template <>
struct Mma16816<__nv_bfloat16> {
  MVB_DEV static void run(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  MVB_DEV static uint32_t pack(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};
template <>
struct Mma16816<__half> {
  MVB_DEV static void run(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, "
        "{%8,%9}, {%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
  MVB_DEV static uint32_t pack(float lo, float hi) {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
};

// ---- scalar conversions for the generic kernels ----
MVB_DEV float to_f32(float x) { return x; }
MVB_DEV float to_f32(__half x) { return __half2float(x); }
MVB_DEV float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
template <typename T>
MVB_DEV T from_f32(float x);
template <>
MVB_DEV float from_f32<float>(float x) { return x; }
template <>
MVB_DEV __half from_f32<__half>(float x) { return __float2half_rn(x); }
template <>
MVB_DEV __nv_bfloat16 from_f32<__nv_bfloat16>(float x) { return __float2bfloat16_rn(x); }

MVB_DEV float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr float kLog2e = 1.4426950408889634f;

}  // namespace mvb200
